"""Multi-GPU sharding of one DGEMM call: column panels of B and C per rank, A cut into K-slabs that are
exchanged rank-to-rank (NVLink) while the GEMMs on already-present slabs run.

The reference's only parallel decomposition is kernel19's row bands over OpenMP threads with a shared packed B
(reference include/kernel19.h:636-656 partition_m_dim, :700, :718-745).  Across GPUs the roles are swapped —
COLUMN panels, A shared — because column panels of column-major B and C are contiguous and no partial sums ever
cross a rank:  C[:, J_r] = beta*C[:, J_r] + alpha * sum_g A[:, K_g] * B[K_g, J_r].

This module is pure host logic (no torch, no CUDA): bench.py drives it with NCCL broadcasts and the CUDA DGEMM,
the CPU tests drive it with gloo broadcasts and a stand-in GEMM.
"""


def partition(total, parts, align):
    """Boundaries [b_0=0, ..., b_parts=total]; every piece is a multiple of `align` except possibly the last
    non-empty one; trailing pieces may be empty when total is small."""
    if parts < 1:
        raise ValueError("parts must be >= 1")
    per = -(-total // parts)          # ceil(total / parts)
    per = -(-per // align) * align    # rounded up to the tile
    return [min(total, i * per) for i in range(parts + 1)]


def slab_order(rank, world):
    """Own slab first (already resident), then ring order — every rank starts on a different slab."""
    return [(rank + i) % world for i in range(world)]


def sharded_dgemm_step(rank, world, kb, beta, start_exchange, wait_exchange, gemm_slab):
    """One DGEMM pass on this rank's column panel.

    start_exchange(g) -> handle   begin making K-slab g (owned by rank g) available locally; called for every g in
                                  the SAME order on every rank (collective-safe)
    wait_exchange(handle)         make the compute stream wait for that slab
    gemm_slab(k_begin, k_count, beta_eff)   C_panel = beta_eff*C_panel + alpha*A[:, k_begin:k_begin+k_count]*B_panel[k_begin:..., :]
    beta is applied exactly once (first slab); later slabs accumulate with beta_eff = 1.
    """
    handles = [start_exchange(g) for g in range(world)]
    first = True
    for g in slab_order(rank, world):
        kc = kb[g + 1] - kb[g]
        if kc <= 0:
            continue
        if g != rank:
            wait_exchange(handles[g])
        gemm_slab(kb[g], kc, beta if first else 1.0)
        first = False
    if first:  # K == 0: the call degenerates to C = beta*C (reference kernel18.h:644-645 order)
        gemm_slab(0, 0, beta)
    wait_exchange(handles[rank])


def sharded_dgemm_step_fused(rank, world, kb, beta, gemm_slab):
    """Same pass, exchange fused INTO the GEMM: the call that multiplies slab i also pulls slab i+1 from its
    owner's memory (dgemm_b200_device_fetch), so no collective and no side stream is needed.

    gemm_slab(k_begin, k_count, beta_eff, fetch)   fetch is None or (owner_rank, k_begin, k_count) of the slab to pull
                                                   from owner_rank's copy of A while this GEMM runs
    A slab is only ever multiplied after the call that fetched it has been issued (stream order makes it complete).
    """
    order = [g for g in slab_order(rank, world) if kb[g + 1] > kb[g]]
    if not order:  # K == 0
        gemm_slab(0, 0, beta, None)
        return
    applied_beta = False
    if order[0] != rank:
        # the own slab is empty (K smaller than the rank count allows): nothing local to start on, so a K = 0 call
        # applies beta and fetches the first slab (no GEMM kernel to hide it behind: the copy engine does it)
        g = order[0]
        gemm_slab(0, 0, beta, (g, kb[g], kb[g + 1] - kb[g]))
        applied_beta = True
    for i, g in enumerate(order):
        nxt = order[i + 1] if i + 1 < len(order) else None
        fetch = (nxt, kb[nxt], kb[nxt + 1] - kb[nxt]) if nxt is not None else None
        gemm_slab(kb[g], kb[g + 1] - kb[g], 1.0 if applied_beta else beta, fetch)
        applied_beta = True
