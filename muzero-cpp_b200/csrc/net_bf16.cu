// net_bf16.cu — placeholder until the tcgen05 towers land (next milestone).
#include "net_backend.h"
namespace mzb {
NetBackend* make_bf16_backend() { return nullptr; }
}
