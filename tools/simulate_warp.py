"""Offline study of warp scheduling policies for the traversal kernels, driven by per-ray step traces from the
CPU oracle (TEST/DEV TOOL; not part of the product). Usage: python tools/simulate_warp.py [n_rays]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import oracle
from gpupathtracer_b200 import Scene, scenes, build_bvh_host

A_INNER, B_TRI, C_ITER = 48, 75, 6      # issue slots per inner step / triangle test / loop bookkeeping


def sequences(ops, off):
    seqs = []
    for i in range(len(off) - 1):
        seqs.append(ops[off[i]:off[i + 1]])
    return seqs


def sim_packet_ifif(seqs):
    """The shader's loop: per iteration [inner block][whole-leaf block], 32-ray packets, no refill."""
    issue = 0; useful = 0
    for w in range(0, len(seqs), 32):
        rays = [list(s) for s in seqs[w:w + 32]]
        pos = [0] * len(rays)
        while True:
            alive = [k for k in range(len(rays)) if pos[k] < len(rays[k])]
            if not alive: break
            issue += C_ITER
            inner = [k for k in alive if rays[k][pos[k]] == 0]
            if inner:
                issue += A_INNER; useful += A_INNER * len(inner)
                for k in inner: pos[k] += 1
            leaf = [k for k in alive if pos[k] < len(rays[k]) and rays[k][pos[k]] == 2]
            if leaf:
                mx = 0
                for k in leaf:
                    pos[k] += 1; c = 0
                    while pos[k] < len(rays[k]) and rays[k][pos[k]] == 1: pos[k] += 1; c += 1
                    useful += B_TRI * c; mx = max(mx, c)
                issue += B_TRI * mx
    return useful / (32.0 * issue), issue


def sim_sched(seqs, leaf_thresh, fetch_thresh, over=14):
    """Voted phases, one triangle per leaf step, dynamic refill; one persistent warp consuming all rays in order
    (models a warp of the persistent grid)."""
    issue = 0; useful = 0
    nxt = 0; n = len(seqs)
    rays = [None] * 32; pos = [0] * 32
    def cur(k):
        r = rays[k]
        while r is not None and pos[k] < len(r) and r[pos[k]] == 2: pos[k] += 1
        if r is None or pos[k] >= len(r): return -1
        return r[pos[k]]
    while True:
        idle = [k for k in range(32) if rays[k] is None or pos[k] >= len(rays[k])]
        if nxt < n and (len(idle) == 32 or len(idle) >= fetch_thresh):
            issue += 30
            for k in idle:
                if nxt < n: rays[k] = seqs[nxt]; pos[k] = 0; nxt += 1
        st = [cur(k) for k in range(32)]
        leaf = [k for k in range(32) if st[k] == 1]; inner = [k for k in range(32) if st[k] == 0]
        if not leaf and not inner:
            if nxt >= n: break
            continue
        issue += over
        if inner and len(leaf) < leaf_thresh:
            issue += A_INNER; useful += A_INNER * len(inner)
            for k in inner: pos[k] += 1
        else:
            issue += B_TRI; useful += B_TRI * len(leaf)
            for k in leaf: pos[k] += 1
    return useful / (32.0 * issue), issue


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
    s = Scene(scenes.atrium())
    nodes, refs, _ = build_bvh_host(s.tris); s.set_bvh(nodes, refs)
    o = oracle.Oracle(s)
    lo, hi = s.bounds()
    for kind, any_hit in (("incoherent closest", False), ("incoherent any", True)):
        rays = scenes.random_rays(n, lo + 0.3, hi - 0.3, seed=5, any_hit=any_hit)
        ops, off = o.trace_ops(rays, any_hit)
        seqs = sequences(ops, off)
        lens = np.diff(off)
        print(f"{kind}: steps/ray mean {lens.mean():.1f} p50 {np.median(lens):.0f} p90 {np.percentile(lens, 90):.0f} max {lens.max()}; "
              f"inner {np.mean(ops == 0) * len(ops) / n:.1f} tris {np.mean(ops == 1) * len(ops) / n:.2f} leaves {np.mean(ops == 2) * len(ops) / n:.2f}")
        e, i0 = sim_packet_ifif(seqs)
        print(f"  packet if-if             : SIMD eff {e:.3f} issue/ray {i0 / n:.0f}")
        for lt in (1, 4, 8, 16):
            for ft in (8, 16, 32):
                e, i1 = sim_sched(seqs, lt, ft)
                print(f"  sched leaf>={lt:2d} fetch>={ft:2d}: SIMD eff {e:.3f} issue/ray {i1 / n:.0f}  ({i0 / i1:.2f}x)")


if __name__ == "__main__":
    main()
