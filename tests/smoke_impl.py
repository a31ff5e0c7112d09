"""smoke(): one small call per hot-path leg on cuda:0, compared with the CPU oracle."""
import numpy as np

import uvmtier
from uvmtier_testlib import gen_prefetch_blocks, load_oracle, run_prefetch_get_hint


def run():
    orc = load_oracle()
    with uvmtier.Context(0) as ctx:
        before = uvmtier.kernel_launch_count()
        inp = gen_prefetch_blocks(256, 3, full_blocks=False, aligned=False, thrashing=True)
        got = ctx.perf_prefetch_get_hint_va_block(inp["blocks"], inp["resident"], inp["faulted"], inp["thrashing"],
                                                  inp["cpu_read"], inp["info"])
        want = run_prefetch_get_hint(orc, "oracle", inp)
        for g, w, name in zip(got, want, ("prefetch mask", "hint residency", "prefetch_info")):
            assert np.array_equal(g, w), "smoke: %s differs from the oracle" % name
        assert uvmtier.kernel_launch_count() > before, "smoke: no CUDA kernel was launched"
    print("smoke ok: %s, %d kernel launches" % (uvmtier.lib().uvmtier_version().decode(), uvmtier.kernel_launch_count() - before))
