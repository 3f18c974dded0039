// include/compat/muzero-cpp/config.h — stands where the reference's muzero-cpp/config.h is included (e.g. tests/mcts_test.cpp,
// self_play.cpp): compile with -I<repo>/include/compat and the reference's names resolve to the B200 engine's drop-in header.
#pragma once
#include "../../muzero_cpp_mzb200.hpp"
