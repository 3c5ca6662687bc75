"""Bit identities the seeding kernel relies on (smem3.cu), checked exhaustively / at random in plain Python."""
import random


def brev32(x):
    return int(f"{x & 0xffffffff:032b}"[::-1], 2)


def keep_loop(vm, cm, l1):
    """The definition: entry L (a bit of vm, all below l1) is kept when cm has a bit in [L, next longer entry)."""
    out, lprev = 0, l1
    for lb in range(31, -1, -1):
        if vm >> lb & 1:
            if cm & ((1 << lprev) - (1 << lb)):
                out |= 1 << lb
            lprev = lb
    return out


def keep_carry(vm, cm, l1):
    """The kernel's branch-free form: bit-reversed, an entry sits on top of its range; adding all-ones over the non-entry
    positions carries a change bit up into the entry's position, where the carry stops."""
    rv = brev32(vm)
    rc = brev32(cm & ((1 << l1) - 1))
    ns = ~rv & 0xffffffff
    return brev32(((((rc & ns) + ns) & 0xffffffff) & rv) | (rc & rv))


def test_group_decision_carry_trick_small_exhaustive():
    for l1 in range(1, 9):
        for vm in range(1 << l1):
            if vm & 1:                                       # entries have L >= 1
                continue
            for cm in range(1 << 9):
                assert keep_carry(vm, cm, l1) == keep_loop(vm, cm, l1), (vm, cm, l1)


def test_group_decision_carry_trick_random():
    rng = random.Random(7)
    for _ in range(200000):
        l1 = rng.randint(1, 15)
        vm = rng.getrandbits(l1) & ~1
        cm = rng.getrandbits(16)
        assert keep_carry(vm, cm, l1) == keep_loop(vm, cm, l1), (vm, cm, l1)


def test_sa40_pack_and_lookup_arithmetic():
    """The word arithmetic of pack_sa40_kernel (prims.cu) and of the packed lookup in dev_sa (fmindex.cuh), restated with
    numpy on random 40-bit values: every row comes back, at all four alignments, with bits 32-39 in use (which no test
    genome reaches: only an index of 2^32 or more rows has such positions)."""
    import numpy as np
    rng = np.random.default_rng(11)
    n = 4099
    v = rng.integers(0, 1 << 40, n, dtype=np.uint64)
    pad = np.zeros((n + 3) // 4 * 4, dtype=np.uint64)
    pad[:n] = v
    q = pad.reshape(-1, 4)
    m32 = np.uint64(0xffffffff)
    w = np.empty((len(q), 5), dtype=np.uint64)
    w[:, 0] = q[:, 0] & m32
    w[:, 1] = ((q[:, 0] >> np.uint64(32)) | (q[:, 1] << np.uint64(8))) & m32
    w[:, 2] = ((q[:, 1] >> np.uint64(24)) | (q[:, 2] << np.uint64(16))) & m32
    w[:, 3] = ((q[:, 2] >> np.uint64(16)) | (q[:, 3] << np.uint64(24))) & m32
    w[:, 4] = (q[:, 3] >> np.uint64(8)) & m32
    words = np.concatenate([w.reshape(-1), np.zeros(2, dtype=np.uint64)])
    assert words.astype(np.uint32).tobytes()[: 5 * n] == b"".join(int(x).to_bytes(5, "little") for x in v)   # = 5 little-endian bytes per row
    k = np.arange(n, dtype=np.uint64)
    b = k * np.uint64(5)
    wi = (b >> np.uint64(2)).astype(np.int64)
    sh = (b & np.uint64(3)) * np.uint64(8)
    got = (((words[wi + 1] << np.uint64(32)) | words[wi]) >> sh) & np.uint64(0xffffffffff)
    assert np.array_equal(got, v)
