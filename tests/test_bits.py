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
