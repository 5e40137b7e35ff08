"""Seeded synthetic DBoW2-style vocabularies (the real ORBvoc.txt is not in the reference checkout) and a second,
pure-Python restatement of TemplatedVocabulary::transform for cross-checking the C++ oracle."""
import numpy as np


def synthetic_vocabulary(k, L, seed, stop_fraction=0.03, ragged=False):
    """k-ary tree of depth L in loadFromTextFile order (breadth-first here; any parent-before-child order is legal):
    children = parent's descriptor with ~25 % of the bits redrawn, leaf weights = idf-like positive doubles, a few 0."""
    rng = np.random.default_rng(seed)
    parent, is_leaf, desc, weight = [], [], [], []
    level = [(0, rng.integers(0, 256, 32, dtype=np.uint8))]
    for depth in range(1, L + 1):
        nxt = []
        for pid, pd in level:
            nk = k if not ragged else int(rng.integers(2, k + 1))
            for _ in range(nk):
                mask = rng.integers(0, 256, 32, dtype=np.uint8) & rng.integers(0, 256, 32, dtype=np.uint8)
                d = pd ^ (mask & rng.integers(0, 256, 32, dtype=np.uint8))
                if rng.random() < 0.02:
                    d = pd.copy()  # identical siblings/parents: distance ties (first child wins)
                parent.append(pid)
                leaf = depth == L
                is_leaf.append(1 if leaf else 0)
                desc.append(d)
                weight.append(0.0 if (leaf and rng.random() < stop_fraction) else float(rng.uniform(0.1, 12.0)) if leaf else 0.0)
                nxt.append((len(parent), d))
        level = nxt
    return np.array(parent, np.int32), np.array(is_leaf, np.uint8), np.stack(desc), np.array(weight, np.float64)


def features_near_words(desc, is_leaf, n, seed):
    """n query descriptors: leaf descriptors with a few flipped bits (so words repeat) plus pure noise."""
    rng = np.random.default_rng(seed)
    leaves = np.flatnonzero(is_leaf)
    pick = rng.choice(leaves, n)
    pick[: n // 5] = pick[n // 5: 2 * (n // 5)]  # repeated words -> addWeight accumulates
    f = desc[pick].copy()
    flips = rng.integers(0, 30, n)
    for i in range(n):
        for b in rng.choice(256, flips[i], replace=False):
            f[i, b >> 3] ^= np.uint8(1 << (b & 7))
    if n >= 10:
        f[n - n // 10:] = rng.integers(0, 256, (n // 10, 32), dtype=np.uint8)
    return f[rng.permutation(n)]


def py_transform(L, parent, is_leaf, desc, weight, feats, levelsup):
    n = len(parent)
    children = [[] for _ in range(n + 1)]
    word_id = [0] * (n + 1)
    nw = 0
    for i in range(n):
        children[parent[i]].append(i + 1)
        if is_leaf[i]:
            word_id[i + 1] = nw
            nw += 1
    bits = np.unpackbits(desc, axis=1)
    bow, fv = {}, {}
    for i, f in enumerate(feats):
        fb = np.unpackbits(f)
        node, lvl, nid = 0, 0, 0
        while True:
            lvl += 1
            ch = children[node]
            d = [(int((bits[c - 1] != fb).sum())) for c in ch]
            node = ch[int(np.argmin(d))]  # first minimum
            if lvl == L - levelsup:
                nid = node
            if not children[node]:
                break
        w = weight[node - 1]
        if w > 0:
            wid = word_id[node]
            bow[wid] = bow[wid] + float(w) if wid in bow else float(w)
            fv.setdefault(nid, []).append(i)
    ids = sorted(bow)
    norm = 0.0
    for k in ids:
        norm += abs(bow[k])
    vals = [bow[k] / norm if norm > 0 else bow[k] for k in ids]
    return ids, vals, {k: fv[k] for k in sorted(fv)}
