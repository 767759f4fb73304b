"""Host-side model of the GEMM chain's tile list (csrc/gemm_chain.cu): static round-robin over CTA pairs, in-order per pair,
a tile starts when its pair is free and its dependencies were published.  Durations in us (1.56 GHz): main loop 0.33 us per
k-block, epilogue + publication 3.5 us (f32 tiles) / 2.5 us (bf16 tiles), 1 us for a counter to be seen.  Prints the makespan
for a few (row blocks, dist) pairs; used to pick FA_CHAIN_DIST.  usage: python profiles/scripts/chain_sim.py"""
import sys

PROBS = [dict(tn=2, kb=8, kind=0), dict(tn=8, kb=8, kind=1), dict(tn=2, kb=32, kind=0), dict(tn=6, kb=8, kind=1)]
KB_US, EPI = 0.33, {0: 3.5, 1: 2.5}
SEE = 1.0


def slots():
    left = [p["tn"] for p in PROBS]
    out = []
    while sum(left):
        best = max((i for i in range(4) if left[i]), key=lambda i: left[i] / PROBS[i]["tn"] + 1e-3 * PROBS[i]["kb"] * 64)
        out.append((best, PROBS[best]["tn"] - left[best]))
        left[best] -= 1
    return out


def simulate(n_rb, dist, pairs=74):
    sl = slots()
    ns = len(sl)
    n_steps = n_rb + 3 * dist
    free = [0.0] * pairs        # when the pair's MMA pipe is free (epilogues overlap the next main loop)
    epi_free = [0.0] * pairs
    pub = {}                    # (prob, rb) -> list of publication times
    for g in range(n_steps * ns):
        t, s = divmod(g, ns)
        pi, tn = sl[s]
        rb = t - pi * dist
        if rb < 0 or rb >= n_rb:
            continue
        c = g % pairs
        ready = 0.0
        if pi > 0:
            d = pub.get((pi - 1, rb), [])
            assert len(d) == PROBS[pi - 1]["tn"], "dependency later in the list"
            ready = max(d) + SEE
        start = max(free[c], ready)
        mma_done = start + PROBS[pi]["kb"] * KB_US
        free[c] = mma_done
        e_start = max(mma_done, epi_free[c])
        e_done = e_start + EPI[PROBS[pi]["kind"]]
        epi_free[c] = e_done
        pub.setdefault((pi, rb), []).append(e_done)
    work = n_rb * sum(p["tn"] * p["kb"] for p in PROBS) * KB_US / pairs
    return max(epi_free), work


for n_rb in (1, 4, 11, 32, 64, 118, 235):
    row = []
    for dist in (1, 2, 4, 9, 16, 24, 32, 48, n_rb):
        if dist > n_rb:
            continue
        mk, work = simulate(n_rb, dist)
        row.append(f"d{dist}:{mk:6.1f}")
    print(f"n_rb {n_rb:4d} (ideal {work:6.1f} us): " + "  ".join(row))
