"""Synthetic replayable-fault traces, block state and page lists for the harness.

All randomness comes from the reference's own test generator (uvm_test_rng, UVM/uvm_test_rng.c:37-62), computed here
in numpy (rng_stream; the library's uvmtier_test_rng_fill is the same stream in C for C callers): one u32 stream per
seed, shaped with numpy.  Nothing in this module loads a shared library, so the CPU reference arm of bench.py builds
its inputs without mapping the product.
Packets are raw 32-byte NVC369_BUF_ENTRY records (UVM/clc369.h:33-66).
"""
import numpy as np

# HW access type encodings, hwref/volta/gv100/dev_fault.h:246-251
HW_VIRT_READ, HW_VIRT_WRITE, HW_VIRT_ATOMIC_STRONG, HW_VIRT_PREFETCH, HW_VIRT_ATOMIC_WEAK = 0, 1, 2, 3, 4
HW_FAULT_TYPE_PTE = 2            # hwref/blackwell/gb100/dev_fault.h:272
HW_FAULT_TYPE_PDE_SIZE = 1       # fatal class (UVM_FAULT_TYPE_PDE_SIZE)
ENG_ID_GRAPHICS = 384            # hwref/blackwell/gb100/dev_fault.h:29
APERTURE_VID = 10                # uvm_aperture_t UVM_APERTURE_VID (UVM/uvm_hal_types.h:33-51)
VA_BASE = 0x7F0000000000
# GPC client ids whose uTLB exists with 10 TPCs per GPC: T1_0..T1_19 (dev_fault.h:287-333)
T1_CLIENTS = np.array(list(range(0x00, 0x08)) + list(range(0x21, 0x29)) + list(range(0x35, 0x39)), np.uint32)

INSTANCE_PTR_DT = np.dtype([("instance_ptr", "<u8"), ("aperture", "u1"), ("ve_id", "u1"), ("va_space", "<u2"),
                            ("gpu_id", "<u2"), ("sort_rank", "<u2")])
assert INSTANCE_PTR_DT.itemsize == 16
GPU_PARAMS_DT = np.dtype([("max_batch_size", "<u4"), ("utlb_per_gpc_count", "<u4"), ("gpc_count", "<u4"),
                          ("va_bits", "<u4"), ("coalesce", "<u4"), ("reserved", "<u4", 3)])
assert GPU_PARAMS_DT.itemsize == 32


_MZ, _MW = 36969 * 65536 - 1, 18000 * 65536 - 1     # moduli of the two multiply-with-carry generators
_LANES = 1 << 16


def _xs_step(x):
    """jsr ^= jsr << 17; jsr ^= jsr >> 13; jsr ^= jsr << 5 on uint32 (scalar or array)."""
    x = x ^ ((x << np.uint32(17)) & np.uint32(0xFFFFFFFF))
    x = x ^ (x >> np.uint32(13))
    return x ^ ((x << np.uint32(5)) & np.uint32(0xFFFFFFFF))


def _xs_apply(cols, x):
    """Applies the GF(2)-linear map given by its 32 columns (images of the unit vectors) to every element of x."""
    y = np.zeros_like(x)
    for i in range(32):
        y ^= np.where((x >> np.uint32(i)) & np.uint32(1), cols[i], np.uint32(0)).astype(np.uint32)
    return y


def _lane_starts(n_lanes, x0, jump):
    """x0, jump(x0), jump(jump(x0)), ... for n_lanes lanes, by doubling: jump is a function (array, k) -> array that
    advances every element by k lane strides."""
    out = np.array([x0], dtype=np.uint64)
    while len(out) < n_lanes:
        out = np.concatenate([out, jump(out, len(out))])
    return out[:n_lanes]


def rng_stream(seed, n):
    """The first n outputs of uvm_test_rng_32 for `seed` (UVM/uvm_test_rng.c:37-62: Marsaglia's KISS with z = 362436069,
    w = 521288629, jcong = 380116160, jsr = seed), in numpy.  The four generators are recurrences; each is run on 65536
    lanes in lockstep, lane j starting at element j * S of the sequence, which jump-ahead makes cheap: a lag-1
    multiply-with-carry state advances by a modular multiplication (state' = state / 2^16 mod a * 2^16 - 1), the
    congruential one by an affine map, the xorshift by a 32 x 32 matrix over GF(2).  Bit-identical to the C loop
    (tests/test_oracle_faults.py compares them)."""
    n = int(n)
    if n <= 0:
        return np.zeros(0, np.uint32)
    R = min(_LANES, n)
    S = (n + R - 1) // R                                   # elements per lane
    with np.errstate(over="ignore"):
        # ---- lane starts ----
        def mwc_starts(s0, m):
            binv_s = pow(pow(65536, -1, m), S, m)          # one lane stride
            def jump(x, k):
                return (x * np.uint64(pow(binv_s, k, m))) % np.uint64(m)
            return _lane_starts(R, s0, jump)
        z = mwc_starts(362436069, _MZ)
        w = mwc_starts(521288629, _MW)
        # x' = 69069 x + 1234567 (mod 2^32): S steps = x -> A x + C
        A, C, a, c, k = 1, 0, 69069, 1234567, S
        while k:
            if k & 1:
                A, C = (A * a) & 0xFFFFFFFF, (C * a + c) & 0xFFFFFFFF
            a, c = (a * a) & 0xFFFFFFFF, (c * a + c) & 0xFFFFFFFF
            k >>= 1
        def lcg_jump(x, kk):
            AA, CC, aa, cc = 1, 0, A, C
            while kk:
                if kk & 1:
                    AA, CC = (AA * aa) & 0xFFFFFFFF, (CC * aa + cc) & 0xFFFFFFFF
                aa, cc = (aa * aa) & 0xFFFFFFFF, (cc * aa + cc) & 0xFFFFFFFF
                kk >>= 1
            return (x * np.uint64(AA) + np.uint64(CC)) & np.uint64(0xFFFFFFFF)
        jc = _lane_starts(R, 380116160, lcg_jump)
        # xorshift: columns of M^S by square-and-multiply on the matrix
        def mat_mul(p, q):                                 # columns of p o q
            return _xs_apply(p, q)
        ident = (np.uint32(1) << np.arange(32, dtype=np.uint32)).astype(np.uint32)
        base = _xs_step(ident.copy())
        ms, bk, k = ident.copy(), base, S
        while k:
            if k & 1:
                ms = mat_mul(bk, ms)
            bk = mat_mul(bk, bk)
            k >>= 1
        def xs_jump(x, kk):
            cols, b2 = ident.copy(), ms
            while kk:
                if kk & 1:
                    cols = mat_mul(b2, cols)
                b2 = mat_mul(b2, b2)
                kk >>= 1
            return _xs_apply(cols, x.astype(np.uint32)).astype(np.uint64)
        js = _lane_starts(R, seed & 0xFFFFFFFF, xs_jump).astype(np.uint32)
        # ---- lockstep ----
        z = z.astype(np.uint64); w = w.astype(np.uint64)
        jc = jc.astype(np.uint32)
        out = np.empty((R, S), np.uint32)
        for t in range(S):
            z = np.uint64(36969) * (z & np.uint64(65535)) + (z >> np.uint64(16))
            w = np.uint64(18000) * (w & np.uint64(65535)) + (w >> np.uint64(16))
            jc = jc * np.uint32(69069) + np.uint32(1234567)
            js = _xs_step(js)
            mwc = (((z & np.uint64(0xFFFF)) << np.uint64(16)) + w).astype(np.uint32)      # (z << 16) + w, 32 bits
            out[:, t] = (mwc ^ jc) + js
    return out.reshape(-1)[:n].copy()


class Stream:
    """Sequential consumer of the KISS u32 stream."""

    def __init__(self, seed, n):
        self.buf = rng_stream(seed, n)
        self.pos = 0

    def take(self, n):
        assert self.pos + n <= len(self.buf), "rng stream exhausted"
        v = self.buf[self.pos:self.pos + n]
        self.pos += n
        return v

    def below(self, n, bound):
        return (self.take(n).astype(np.uint64) % np.uint64(bound)).astype(np.int64)

    def unit(self, n):
        return self.take(n).astype(np.float64) / 4294967296.0


def gpu_params(max_batch_size=256, tpc_per_gpc=10, gpc_count=8, va_bits=57, coalesce=1):
    """uvm_parent_gpu_t fields the fault path reads (Blackwell: utlb_per_gpc = 2*TPC/GPC + 1,
    UVM/uvm_blackwell_fault_buffer.h:85-90)."""
    g = np.zeros(1, GPU_PARAMS_DT)
    g["max_batch_size"] = max_batch_size
    g["utlb_per_gpc_count"] = 2 * tpc_per_gpc + 1
    g["gpc_count"] = gpc_count
    g["va_bits"] = va_bits
    g["coalesce"] = coalesce
    return g


def instance_ptr_table(n_spaces, gpu_id=1):
    """One instance pointer (channel) per VA space, all on `gpu_id`."""
    t = np.zeros(n_spaces, INSTANCE_PTR_DT)
    t["instance_ptr"] = (0x1000 + np.arange(n_spaces, dtype=np.uint64) * 7) << np.uint64(12)
    t["aperture"] = APERTURE_VID
    t["ve_id"] = 0xFF
    t["va_space"] = np.arange(n_spaces)
    t["gpu_id"] = gpu_id
    t["sort_rank"] = np.arange(n_spaces)     # (va_space, gpu) pairs are already in order
    return t


def encode_packets(inst_ptr, address, access_hw, gpc, client, timestamp, veid=None, fault_type=None, valid=None,
                   inst_aperture_hw=None):
    """Vectorised NVC369_BUF_ENTRY encoder (UVM/clc369.h:33-66). address/inst_ptr in bytes (4K aligned)."""
    n = len(address)
    w = np.zeros((n, 8), np.uint32)
    ip = np.asarray(inst_ptr, np.uint64) >> np.uint64(12)
    ap = np.zeros(n, np.uint32) if inst_aperture_hw is None else np.asarray(inst_aperture_hw, np.uint32)
    w[:, 0] = ((ip & np.uint64(0xFFFFF)) << np.uint64(12)).astype(np.uint32) | (ap << 8)
    w[:, 1] = (ip >> np.uint64(20)).astype(np.uint32)
    pg = (np.asarray(address, np.uint64) >> np.uint64(12)) & np.uint64((1 << 52) - 1)
    w[:, 2] = ((pg & np.uint64(0xFFFFF)) << np.uint64(12)).astype(np.uint32)
    w[:, 3] = (pg >> np.uint64(20)).astype(np.uint32)
    ts = np.asarray(timestamp, np.uint64)
    w[:, 4] = (ts & np.uint64(0xFFFFFFFF)).astype(np.uint32)
    w[:, 5] = (ts >> np.uint64(32)).astype(np.uint32)
    ve = np.zeros(n, np.uint32) if veid is None else np.asarray(veid, np.uint32)
    w[:, 6] = ENG_ID_GRAPHICS + ve
    ft = np.full(n, HW_FAULT_TYPE_PTE, np.uint32) if fault_type is None else np.asarray(fault_type, np.uint32)
    vd = np.ones(n, np.uint32) if valid is None else np.asarray(valid, np.uint32)
    w[:, 7] = (ft | (1 << 7) | (np.asarray(client, np.uint32) << 8) | (np.asarray(access_hw, np.uint32) << 16) |
               (0 << 20) | (np.asarray(gpc, np.uint32) << 24) | (1 << 30) | (vd << 31))
    return np.ascontiguousarray(w)


def fault_trace(n, seed=1, distribution="uniform", n_spaces=4, total_va_bytes=64 << 30, gpc_count=8,
                gpu_id=1):
    """Returns (packets[n,8] u32, table, layout) for one GPU's ring.

    Access mix 60/25/5/8/2 % READ/WRITE/ATOMIC_STRONG/PREFETCH/ATOMIC_WEAK.  The VA of every VA space
    starts at VA_BASE and spans total_va_bytes / n_spaces (2 MB aligned).
    distribution: uniform | stream | hotset | adversarial
    """
    s = Stream(seed, 12 * n + 64)
    blocks_per_space = max(1, (total_va_bytes >> 21) // n_spaces)
    pages_per_space = blocks_per_space * 512
    table = instance_ptr_table(n_spaces, gpu_id)
    space = s.below(n, n_spaces)
    gpc = s.below(n, gpc_count)
    client = T1_CLIENTS[s.below(n, len(T1_CLIENTS))]
    u = s.unit(n)
    access = np.select([u < 0.60, u < 0.85, u < 0.90, u < 0.98], [HW_VIRT_READ, HW_VIRT_WRITE, HW_VIRT_ATOMIC_STRONG,
                                                                 HW_VIRT_PREFETCH], HW_VIRT_ATOMIC_WEAK)
    if distribution == "uniform":
        page = s.below(n, pages_per_space)
    elif distribution == "stream":
        # sequential pages, each touched d in [1,8] times (different uTLBs come from gpc/client draws)
        d = 1 + s.below(n, 8)
        reps = np.repeat(np.arange(n), d)[:n]
        page = reps % pages_per_space
        space = (reps // pages_per_space) % n_spaces
    elif distribution == "hotset":
        hot_blocks = max(1, blocks_per_space // 100)
        hot = s.unit(n) < 0.9
        blk = np.where(hot, s.below(n, hot_blocks) * (blocks_per_space // hot_blocks), s.below(n, blocks_per_space))
        page = blk * 512 + s.below(n, 512)
    elif distribution == "adversarial":
        # few pages, rising/falling access types, alternating uTLBs: stresses the fetch-merge rules
        page = s.below(n, 6) + 512 * s.below(n, 2)
        gpc = s.below(n, 2)
        client = T1_CLIENTS[s.below(n, 3)]
        space = s.below(n, min(2, n_spaces))
        access = s.below(n, 5)
    else:
        raise ValueError(distribution)
    addr = np.uint64(VA_BASE) + (page.astype(np.uint64) << np.uint64(12))
    inst = table["instance_ptr"][space]
    ts = np.uint64(1000) + np.arange(n, dtype=np.uint64) * np.uint64(3)
    pk = encode_packets(inst, addr, access, gpc, client, ts)
    layout = dict(va_base=VA_BASE, blocks_per_space=blocks_per_space, n_spaces=n_spaces, gpu_id=gpu_id)
    return pk, table, layout
