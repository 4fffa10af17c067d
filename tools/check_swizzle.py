"""Enumerates every fragment LDS.64 of the DGEMM warp tile against the SWIZZLE_128B shared-memory layouts and checks
that it is bank-conflict free (each half-warp of an 8-byte load must touch 16 distinct 8-byte bank pairs).
Mirrors a_smem_off / b_smem_off / perm_n / map_m of csrc/dgemm_device.cuh.  Run: python tools/check_swizzle.py"""
MAP_M = [0, 1, 8, 9, 2, 3, 10, 11]   # mma column c -> m offset inside a 16-row sub-tile (second fragment: +4)
PERM_N = [0, 2, 4, 6, 1, 3, 5, 7]    # mma row r -> n offset inside an 8-column block


def a_off(m_local, k_local):
    mm = m_local & 15
    return (m_local >> 4) * 2048 + k_local * 128 + (((mm >> 1) ^ (k_local & 7)) << 4) + ((mm & 1) << 3)


def b_off(k_local, n_local):
    return n_local * 128 + (((k_local >> 1) ^ (n_local & 7)) << 4) + ((k_local & 1) << 3)


def conflict_free(addrs):
    return all(len({(a % 128) // 8 for a in addrs[16 * h:16 * h + 16]}) == 16 for h in range(2))


def main():
    ok = True
    for kk in range(4):
        for f in range(2):
            for s in range(8):
                ok &= conflict_free([a_off(16 * s + 4 * f + MAP_M[l // 4], kk * 4 + l % 4) for l in range(32)])
        for t in range(16):
            ok &= conflict_free([b_off(kk * 4 + l % 4, 8 * t + PERM_N[l // 4]) for l in range(32)])
    # A split by k parity (dgemm_tma_ws_split_kernel, LaneOffsetsKS): k in row (k&1)*8 + (k>>1); lanes q of k-step kk feed
    # k = 8*(kk>>1) + 2q + ((q^kk)&1); B rows permuted by PERM_N_KS
    PERM_N_KS = [0, 1, 4, 5, 2, 3, 6, 7]
    ks_ok = True
    for kk in range(4):
        ks = [8 * (kk >> 1) + 2 * q + ((q ^ kk) & 1) for q in range(4)]
        for f in range(2):
            for s in range(8):
                ks_ok &= conflict_free([a_off(16 * s + 4 * f + MAP_M[l // 4], (ks[l % 4] & 1) * 8 + (ks[l % 4] >> 1))
                                        for l in range(32)])
        for t in range(16):
            ks_ok &= conflict_free([b_off(ks[l % 4], 8 * t + PERM_N_KS[l // 4]) for l in range(32)])
    assert sorted(sum(([8 * (kk >> 1) + 2 * q + ((q ^ kk) & 1) for q in range(4)] for kk in range(4)), [])) == list(range(16))
    print(f"k-parity-split A layout conflict-free: {ks_ok}")
    assert ks_ok
    natural_a = conflict_free([a_off(l // 4, l % 4) for l in range(32)])
    natural_b = conflict_free([b_off(l % 4, l // 4) for l in range(32)])
    print(f"permuted lane maps conflict-free: {ok}; natural maps: A {natural_a}, B {natural_b}")
    # every (m, k) / (k, n) of a stage maps to a distinct 8-byte slot
    assert len({a_off(m, k) for m in range(128) for k in range(16)}) == 2048
    assert len({b_off(k, n) for k in range(16) for n in range(128)}) == 2048
    assert ok and not natural_a and not natural_b
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
