"""oracle_py — TEST INFRASTRUCTURE.  ctypes bindings for the C search oracle (oracle/mcts_oracle.c), a parser
for `ref_driver` text dumps, and the MZBT tensor container reader/writer.  Only tests/, smoke() and bench.py's
cpu_baseline leg import this."""
import ctypes as C
import os
import struct
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
MAX_A = 32
LIB_PATH = os.path.join(HERE, "_build", "libmzb_oracle.so")
REF_DRIVER = os.path.join(HERE, "_ref", "ref_driver")


def build_oracle(force=False):
    """gcc-compile the C restatement (no FMA contraction, like the reference's -O3 x86-64 build)."""
    src = os.path.join(HERE, "mcts_oracle.c")
    deps = [src, os.path.join(HERE, "mcts_oracle.h"), os.path.join(HERE, "se_evaluator.h")]
    if (not force and os.path.exists(LIB_PATH)
            and all(os.path.getmtime(LIB_PATH) >= os.path.getmtime(d) for d in deps)):
        return LIB_PATH
    os.makedirs(os.path.dirname(LIB_PATH), exist_ok=True)
    subprocess.check_call(["gcc", "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-Wall", "-o", LIB_PATH, src, "-lm"])
    return LIB_PATH


class Config(C.Structure):
    _fields_ = [("num_players", C.c_int32), ("num_simulations", C.c_int32), ("num_actions", C.c_int32),
                ("dirichlet_alpha", C.c_double), ("dirichlet_epsilon", C.c_double), ("pb_c_base", C.c_double),
                ("pb_c_init", C.c_double), ("discount", C.c_double), ("value_lowerbound", C.c_double),
                ("value_upperbound", C.c_double)]


class SEParams(C.Structure):
    _fields_ = [("vlo", C.c_double), ("vhi", C.c_double), ("rlo", C.c_double), ("rhi", C.c_double)]


class Result(C.Structure):
    _fields_ = [("root_value", C.c_double), ("max_tree_depth", C.c_int32), ("n_children", C.c_int32),
                ("child_actions", C.c_int32 * MAX_A), ("child_visits", C.c_int32 * MAX_A),
                ("children_relative_visit", C.c_double * MAX_A), ("emperical_policy", C.c_double * MAX_A)]


class Node(C.Structure):
    _fields_ = [("visit_count", C.c_int32), ("to_play", C.c_int32), ("action", C.c_int32), ("prior", C.c_double),
                ("value_sum", C.c_double), ("reward", C.c_double), ("child_base", C.c_int32),
                ("n_children", C.c_int32), ("node_id", C.c_int32)]


class MinMax(C.Structure):
    _fields_ = [("init_min", C.c_double), ("init_max", C.c_double), ("min", C.c_double), ("max", C.c_double)]


class MT(C.Structure):
    _fields_ = [("mt", C.c_uint32 * 624), ("idx", C.c_int32)]


class MCTS(C.Structure):
    _fields_ = [("cfg", Config), ("minmax", MinMax), ("rng", MT), ("nodes", C.POINTER(Node)), ("n_nodes", C.c_int32)]


INF_D = float(np.finfo(np.float64).max)      # types.h:78
NINF_D = -float(np.finfo(np.float64).max)    # types.h:77 (numeric_limits<double>::lowest())

_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(build_oracle())
        _lib.mzo_mcts_create.restype = C.POINTER(MCTS)
        _lib.mzo_mcts_create.argtypes = [C.POINTER(Config), C.c_int32]
        _lib.mzo_mcts_destroy.argtypes = [C.POINTER(MCTS)]
        _lib.mzo_ucb_score.restype = C.c_double
        _lib.mzo_ucb_score.argtypes = [C.POINTER(MCTS), C.POINTER(Node), C.POINTER(Node)]
        _lib.mzo_select_child.restype = C.c_int
        _lib.mzo_select_child.argtypes = [C.POINTER(MCTS), C.c_int]
        _lib.mzo_backpropagate.argtypes = [C.POINTER(MCTS), C.POINTER(C.c_int32), C.c_int, C.c_double, C.c_int]
        _lib.mzo_expand.argtypes = [C.POINTER(MCTS), C.c_int, C.POINTER(C.c_int32), C.c_int, C.c_int, C.c_double,
                                    C.POINTER(C.c_double), C.c_int]
        _lib.mzo_add_exploration_noise.argtypes = [C.POINTER(MCTS), C.c_int, C.c_double, C.c_double]
        _lib.mzo_minmax_init.argtypes = [C.POINTER(MinMax), C.c_double, C.c_double]
        _lib.mzo_minmax_reset.argtypes = [C.POINTER(MinMax)]
        _lib.mzo_minmax_update.argtypes = [C.POINTER(MinMax), C.c_double]
        _lib.mzo_minmax_normalize.restype = C.c_double
        _lib.mzo_minmax_normalize.argtypes = [C.POINTER(MinMax), C.c_double]
        _lib.mzo_mt_seed.argtypes = [C.POINTER(MT), C.c_uint32]
        _lib.mzo_mt_next.restype = C.c_uint32
        _lib.mzo_mt_next.argtypes = [C.POINTER(MT)]
        _lib.mzo_shuffle_perm.argtypes = [C.POINTER(MT), C.c_int, C.POINTER(C.c_int)]
        _lib.mzo_sample_dirichlet.argtypes = [C.POINTER(MT), C.c_double, C.c_int, C.POINTER(C.c_double)]
        _lib.mzo_run_se.argtypes = [C.POINTER(Config), C.POINTER(SEParams), C.c_int, C.c_int, C.c_int, C.c_int,
                                    C.c_int, C.POINTER(Result), C.POINTER(C.c_int32), C.POINTER(C.c_int32)]
        _lib.mzo_run_replay.argtypes = [C.POINTER(Config), C.c_int, C.c_uint32, C.c_int, C.c_int,
                                        C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_double),
                                        C.POINTER(Result), C.POINTER(C.c_int32), C.POINTER(C.c_int32)]
    return _lib


def make_config(num_players=2, num_simulations=50, num_actions=7, alpha=0.3, eps=0.25, pb_c_base=19652.0,
                pb_c_init=1.25, discount=0.997, lower=-10.0, upper=10.0):
    return Config(num_players, num_simulations, num_actions, alpha, eps, pb_c_base, pb_c_init, discount, lower, upper)


def result_to_dict(r, num_actions, trace_parent=None, trace_action=None):
    n = r.n_children
    d = dict(root_value=r.root_value, depth=r.max_tree_depth, actions=list(r.child_actions[:n]),
             visits=list(r.child_visits[:n]), relvisits=list(r.children_relative_visit[:n]),
             policy=list(r.emperical_policy[:num_actions]))
    if trace_parent is not None:
        d["trace"] = list(zip([int(x) for x in trace_parent], [int(x) for x in trace_action]))
    return d


def run_se(cfg, seed, tree, runs=1, legal_mode=0, noise=True, se=(-1.0, 1.0, -0.5, 0.5)):
    """One tree slot (MCTS object seeded seed+tree), `runs` consecutive searches; mirrors `ref_driver mcts`."""
    res = (Result * runs)()
    n = cfg.num_simulations
    tp = (C.c_int32 * (runs * n))()
    ta = (C.c_int32 * (runs * n))()
    sep = SEParams(*se)
    lib().mzo_run_se(C.byref(cfg), C.byref(sep), seed, tree, runs, legal_mode, int(noise), res, tp, ta)
    return [result_to_dict(res[r], cfg.num_actions, tp[r * n:(r + 1) * n], ta[r * n:(r + 1) * n]) for r in range(runs)]


def run_replay(cfg, seed, legal_mask, to_play, noise, value, reward, policy):
    """Search one tree on recorded per-node network outputs (value/reward [n+1], policy [n+1, A])."""
    value = np.ascontiguousarray(value, dtype=np.float64)
    reward = np.ascontiguousarray(reward, dtype=np.float64)
    policy = np.ascontiguousarray(policy, dtype=np.float64)
    n = cfg.num_simulations
    assert value.shape == (n + 1,) and policy.shape == (n + 1, cfg.num_actions)
    res = Result()
    tp = (C.c_int32 * n)()
    ta = (C.c_int32 * n)()
    dp = C.POINTER(C.c_double)
    lib().mzo_run_replay(C.byref(cfg), int(seed), int(legal_mask), int(to_play), int(noise), value.ctypes.data_as(dp),
                         reward.ctypes.data_as(dp), policy.ctypes.data_as(dp), C.byref(res), tp, ta)
    return result_to_dict(res, cfg.num_actions, tp, ta)


# ---------------------------------------------------------------- ref_driver text dumps
def parse_ref_mcts(text):
    """Parse `ref_driver mcts|replay` output -> list of dicts (same keys as result_to_dict, relvisits as floats)."""
    out = []
    cur = None
    for line in text.splitlines():
        tok = line.split()
        if not tok:
            continue
        if tok[0] == "tree":
            cur = dict(tree=int(tok[1]), run=int(tok[3]), root_value=float.fromhex(tok[5]), depth=int(tok[7]))
            out.append(cur)
        elif tok[0] == "actions":
            cur["actions"] = [int(x) for x in tok[1:]]
        elif tok[0] == "relvisits":
            cur["relvisits"] = [float.fromhex(x) for x in tok[1:]]
        elif tok[0] == "policy":
            cur["policy"] = [float.fromhex(x) for x in tok[1:]]
        elif tok[0] == "trace":
            cur["trace"] = [tuple(int(y) for y in x.split(":")) for x in tok[1:]]
    return out


def have_ref():
    return os.path.exists(REF_DRIVER)


def run_ref(*args):
    return subprocess.run([REF_DRIVER] + [str(a) for a in args], check=True, capture_output=True, text=True).stdout


# ---------------------------------------------------------------- MZBT container (oracle/mzbt_io.h)
_DT = {0: np.float32, 1: np.float64, 2: np.int32, 3: np.uint8}
_DT_INV = {np.dtype(np.float32): 0, np.dtype(np.float64): 1, np.dtype(np.int32): 2, np.dtype(np.uint8): 3}


def mzbt_load(path):
    with open(path, "rb") as f:
        buf = f.read()
    assert buf[:4] == b"MZBT", "not an MZBT file"
    ver, n = struct.unpack_from("<II", buf, 4)
    off = 12
    out = {}
    for _ in range(n):
        (nl,) = struct.unpack_from("<I", buf, off); off += 4
        name = buf[off:off + nl].decode(); off += nl
        dt, nd = struct.unpack_from("<II", buf, off); off += 8
        dims = struct.unpack_from("<%dq" % nd, buf, off); off += 8 * nd
        cnt = int(np.prod(dims)) if nd else 1
        arr = np.frombuffer(buf, dtype=_DT[dt], count=cnt, offset=off).reshape(dims).copy()
        off += cnt * np.dtype(_DT[dt]).itemsize
        out[name] = arr
    return out


def mzbt_save(path, tensors):
    with open(path, "wb") as f:
        f.write(b"MZBT" + struct.pack("<II", 1, len(tensors)))
        for name, arr in tensors.items():
            arr = np.ascontiguousarray(arr)
            nb = name.encode()
            f.write(struct.pack("<I", len(nb)) + nb + struct.pack("<II", _DT_INV[arr.dtype], arr.ndim))
            f.write(struct.pack("<%dq" % arr.ndim, *arr.shape))
            f.write(arr.tobytes())


# ---------------------------------------------------------------- custom evaluators (Python callbacks)
class EvalOut(C.Structure):
    _fields_ = [("value", C.c_double), ("reward", C.c_double), ("policy", C.c_double * MAX_A)]


EVAL_FN = C.CFUNCTYPE(None, C.c_void_p, C.c_int, C.c_int, C.c_int, C.POINTER(EvalOut))


def mcts_create(cfg, seed):
    return lib().mzo_mcts_create(C.byref(cfg), seed)


def mcts_destroy(m):
    lib().mzo_mcts_destroy(m)


def mcts_run(m, legal, to_play, noise, eval_py):
    """eval_py(parent_node, action, new_node) -> (value, reward, policy list); parent_node < 0 = initial inference."""
    l = lib()
    l.mzo_mcts_run.argtypes = [C.POINTER(MCTS), C.POINTER(C.c_int32), C.c_int, C.c_int, C.c_int, EVAL_FN, C.c_void_p,
                               C.POINTER(Result), C.POINTER(C.c_int32), C.POINTER(C.c_int32)]
    cfg = m.contents.cfg

    def cb(ctx, parent, action, new_node, out):
        v, r, p = eval_py(parent, action, new_node)
        out.contents.value = v
        out.contents.reward = r
        for i, x in enumerate(p):
            out.contents.policy[i] = x

    fn = EVAL_FN(cb)
    n = cfg.num_simulations
    res = Result()
    tp = (C.c_int32 * max(n, 1))()
    ta = (C.c_int32 * max(n, 1))()
    arr = (C.c_int32 * len(legal))(*legal)
    l.mzo_mcts_run(m, arr, len(legal), to_play, int(noise), fn, None, C.byref(res), tp, ta)
    return result_to_dict(res, cfg.num_actions, tp[:n], ta[:n])
