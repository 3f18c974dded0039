"""Regenerate the self-play helper fixtures from the reference build (oracle/_ref/ref_driver, modes `stack` and `select`).
Run in the build container only:   make -C oracle/ref_build && python tests/golden/make_golden_selfplay.py
Outputs (committed): selfplay_c4.txt (whole games through the unmodified play_game / reanalyze, see main_c4), selfplay_stack.txt (GameHistory::get_stacked_observations, types.cpp:16-42, on synthetic games) and
selfplay_select.txt (the self_play.cpp:21-39 selection rule over libstdc++'s std::discrete_distribution / std::mt19937)."""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
DRV = os.path.join(ROOT, "oracle", "_ref", "ref_driver")

STACK_CASES = [dict(c=2, h=3, w=4, ach=1, stack=3, steps=6, actions=5), dict(c=3, h=6, w=7, ach=1, stack=0, steps=3, actions=7),
               dict(c=1, h=4, w=4, ach=2, stack=6, steps=9, actions=18), dict(c=3, h=3, w=3, ach=1, stack=2, steps=2, actions=9)]
SELECT_CASES = [dict(seed=0, draws=60, actions=7, temperature=1.0), dict(seed=5, draws=60, actions=7, temperature=0.5),
                dict(seed=9, draws=60, actions=18, temperature=0.25), dict(seed=2, draws=40, actions=9, temperature=0.0),
                dict(seed=11, draws=40, actions=3, temperature=2.0)]


def run(mode, kw):
    return subprocess.run([DRV, mode] + ["%s=%s" % kv for kv in kw.items()], check=True, capture_output=True, text=True).stdout


def main_c4():
    """tests/golden/selfplay_c4.txt: 128 Connect4 games played by the UNMODIFIED play_game (self_play.cpp:84-183), each re-searched
    by the UNMODIFIED reanalyze (self_play.cpp:186-225), plus 16 evaluator games against the Random opponent — all over the
    observation-keyed synthetic evaluator (oracle/ref_build/ref_selfplay.cpp)."""
    import subprocess
    drv = os.path.join(ROOT, "oracle", "_ref", "ref_selfplay")
    subprocess.check_call([drv, "games=128", "sims=50", "eval_games=16", "out=" + os.path.join(HERE, "selfplay_c4.txt"),
                           # replay_c4.txt: the same 128 histories through the unmodified PrioritizedReplayBuffer (capacity 1000 < stored
                           # positions, so games are evicted), a sampled Batch every 8 games — golden for mzb_replay_* (csrc/replay.cu)
                           "replay_out=" + os.path.join(HERE, "replay_c4.txt"), "replay_capacity=1000", "replay_batch=16", "replay_stack=2"])


if __name__ == "__main__":
    main_c4()
    with open(os.path.join(HERE, "selfplay_stack.txt"), "w") as f:
        for kw in STACK_CASES:
            f.write(run("stack", kw))
    with open(os.path.join(HERE, "selfplay_select.txt"), "w") as f:
        for kw in SELECT_CASES:
            f.write(run("select", kw))
    print("written")
