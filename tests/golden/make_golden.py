"""Regenerates the committed golden vectors from the REFERENCE's own C (oracle/_ref/*.so, built
from /root/reference by oracle/build_ref.sh).  Run in the build container only:

    python tests/golden/make_golden.py

The .npz files hold inputs and the reference's outputs; tests compare the oracle restatement,
the SIMT-emulated kernels and the CUDA path against them.  /root/reference is not needed to USE
the fixtures.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import uvmtier_testlib as T  # noqa: E402


def prefetch():
    ref = T.load_ref("prefetch")
    assert ref is not None, "build oracle/_ref first (needs /root/reference)"
    cases = {}
    k = 0
    for cfg in (dict(), dict(full_blocks=False), dict(full_blocks=False, aligned=False),
                dict(full_blocks=False, aligned=False, thrashing=True),
                dict(dst="cpu", full_blocks=False, aligned=False, thrashing=True)):
        for thr in (51, 30, 100):
            inp = T.gen_prefetch_blocks(64, 1000 + k, **cfg)
            mask, res, info = T.run_prefetch_get_hint(ref, "ref", inp, (1, thr, 1))
            pre = "hint%d_" % k
            cases[pre + "threshold"] = np.array(thr)
            for name in ("blocks", "resident", "faulted", "info"):
                cases[pre + name] = inp[name]
            cases[pre + "thrashing"] = inp["thrashing"] if inp["thrashing"] is not None else np.zeros((0, 8), "<u8")
            cases[pre + "cpu_read"] = inp["cpu_read"] if inp["cpu_read"] is not None else np.zeros((0, 8), "<u8")
            cases[pre + "out_mask"] = mask
            cases[pre + "out_residency"] = res
            cases[pre + "out_info"] = info
            k += 1
    cases["n_hint"] = np.array(k)
    k = 0
    for thr in (51, 1):
        for with_res in (True, False):
            inp = T.gen_compute_ats(64, 2000 + k)
            out = T.run_compute_ats(ref, "ref", inp, with_res, (1, thr, 1))
            pre = "ats%d_" % k
            cases[pre + "threshold"] = np.array(thr)
            cases[pre + "with_residency"] = np.array(with_res)
            for name in ("faulted", "residency", "fregion", "mregion"):
                cases[pre + name] = inp[name]
            cases[pre + "out_mask"] = out
            k += 1
    cases["n_ats"] = np.array(k)
    np.savez_compressed(os.path.join(HERE, "prefetch_golden.npz"), **cases)
    print("wrote prefetch_golden.npz", sum(v.nbytes for v in cases.values()), "bytes raw")


def faults():
    import faultlib as F
    from uvmtier import tracegen as TG
    ref = T.load_ref("faults")
    assert ref is not None, "build oracle/_ref first (needs /root/reference)"
    cases = {}
    k = 0
    for dist, ns, B, n in (("uniform", 4, 256, 700), ("stream", 1, 256, 700), ("hotset", 4, 64, 500),
                           ("adversarial", 2, 256, 900), ("adversarial", 4, 100, 600), ("stream", 16, 1000, 1500)):
        pk, tbl, _ = TG.fault_trace(n, 77 + k, dist, ns)
        gp = TG.gpu_params(B)
        r = F.run_fetch_preprocess(ref, "ref", pk, gp, tbl)
        pre = "c%d_" % k
        cases[pre + "packets"] = pk
        cases[pre + "table"] = tbl
        cases[pre + "gpu_params"] = gp
        for name in ("entries", "ordered", "va_space", "gpu", "batch", "ring_after"):
            cases[pre + name] = r[name]
        k += 1
    pk, tbl = F.directed_quirk_trace()
    gp = TG.gpu_params()
    r = F.run_fetch_preprocess(ref, "ref", pk, gp, tbl)
    pre = "c%d_" % k
    cases[pre + "packets"] = pk; cases[pre + "table"] = tbl; cases[pre + "gpu_params"] = gp
    for name in ("entries", "ordered", "va_space", "gpu", "batch", "ring_after"):
        cases[pre + name] = r[name]
    k += 1
    cases["n_cases"] = np.array(k)
    np.savez_compressed(os.path.join(HERE, "faults_golden.npz"), **cases)
    print("wrote faults_golden.npz")


if __name__ == "__main__":
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(HERE)), "open-gpu-kernel-modules_b200"))
    prefetch()
    faults()
