"""numpy restatement of csrc/sampler.cuh (test infrastructure): llama.cpp's default chain [external] —
top-k cut -> nucleus on T=1 probabilities -> temperature -> categorical draw with Philox4x32-10(seed, step)."""
import numpy as np

CAND = 1024


def philox4x32_10(seed, ctr):
    M = 0xFFFFFFFF
    c = [ctr & M, (ctr >> 32) & M, 0, 0]
    k0, k1 = seed & M, (seed >> 32) & M
    for _ in range(10):
        p0, p1 = 0xD2511F53 * c[0], 0xCD9E8D57 * c[2]
        c = [((p1 >> 32) ^ c[1] ^ k0) & M, p1 & M, ((p0 >> 32) ^ c[3] ^ k1) & M, p0 & M]
        k0 = (k0 + 0x9E3779B9) & M
        k1 = (k1 + 0xBB67AE85) & M
    return c


def sample(logits, temperature, top_k=0, top_p=1.0, seed=0, step=0, return_dist=False):
    lg = np.asarray(logits, dtype=np.float32)
    V = len(lg)
    K = min(top_k, V) if 0 < top_k < CAND else min(CAND, V)
    order = np.lexsort((np.arange(V), -lg.astype(np.float64)))[:K]        # logit desc, index asc
    cv = lg[order].astype(np.float64)
    e1 = np.exp(cv - cv[0])
    keep = K
    if top_p < 1.0:
        cum = np.cumsum(e1)
        hit = np.nonzero(cum >= np.float32(top_p) * cum[-1])[0]
        keep = int(hit[0]) + 1 if len(hit) else K
    w = np.exp((cv[:keep] - cv[0]) / np.float32(temperature))
    cum = np.cumsum(w)
    r = philox4x32_10(seed, step)
    u = np.float32(r[0] >> 8) * np.float32(1.0 / 16777216.0) * cum[-1]
    hit = np.nonzero(cum > u)[0]
    pick = int(hit[0]) if len(hit) else keep - 1
    if return_dist:
        return int(order[pick]), order[:keep], w / cum[-1], u / cum[-1]
    return int(order[pick])
