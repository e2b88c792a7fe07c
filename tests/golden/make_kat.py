"""Writes tests/golden/kat_d4.json: hand-checkable known answers for the flat-search
semantics (SURVEY.md section 8c "Known-answer mini-fixtures").  Everything is computed
with exact rational arithmetic (fractions) from the definitions -- squared L2 =
sum (q-x)^2, IP = sum q*x, LangChain relevance = 1 - d/sqrt(2) -- not with the oracle,
so the oracle itself is checked against this file.

Run:  python tests/golden/make_kat.py
"""
import json
import math
import os
from fractions import Fraction as F

HERE = os.path.dirname(os.path.abspath(__file__))


def l2(q, x):
    return sum((F(a) - F(b)) ** 2 for a, b in zip(q, x))


def ip(q, x):
    return sum(F(a) * F(b) for a, b in zip(q, x))


def topk(metric, q, xs, k):
    sc = [(l2(q, x) if metric == "l2" else ip(q, x), i) for i, x in enumerate(xs)]
    sc.sort(key=(lambda t: (t[0], t[1])) if metric == "l2" else (lambda t: (-t[0], t[1])))
    ids = [i for _, i in sc[:k]] + [-1] * max(0, k - len(sc))
    vals = [float(s) for s, _ in sc[:k]] + [None] * max(0, k - len(sc))  # None = +-FLT_MAX padding
    return vals, ids


E = [[1, 0, 0, 0], [0, 1, 0, 0], [0, 0, 1, 0], [0, 0, 0, 1]]
cases = []

# identical vector: d = 0, relevance 1.0; orthogonal unit vectors: d = 2, rel = 1 - sqrt(2) < 0
xs = E
q = [1, 0, 0, 0]
for metric in ("l2", "ip"):
    v, i = topk(metric, q, xs, 4)
    cases.append(dict(name=f"unit_axes_{metric}", metric=metric, corpus=xs, query=q, k=4, scores=v, ids=i))

# k > ntotal: -1 padding
xs = [[1, 2, 3, 4], [0, 0, 0, 0]]
q = [1, 1, 1, 1]
for metric in ("l2", "ip"):
    v, i = topk(metric, q, xs, 5)
    cases.append(dict(name=f"k_gt_ntotal_{metric}", metric=metric, corpus=xs, query=q, k=5, scores=v, ids=i))

# exact ties: lowest id first
xs = [[2, 0, 0, 0], [0, 2, 0, 0], [1, 1, 1, 1], [0, 0, 2, 0], [3, 3, 3, 3]]
q = [0, 0, 0, 0]
v, i = topk("l2", q, xs, 3)
cases.append(dict(name="ties_l2", metric="l2", corpus=xs, query=q, k=3, scores=v, ids=i))
q = [1, 1, 1, 0]
v, i = topk("ip", q, xs, 4)
cases.append(dict(name="ties_ip", metric="ip", corpus=xs, query=q, k=4, scores=v, ids=i))

# quarter-integers (exact in fp32), negative coordinates
xs = [[0.25, -0.5, 0.75, 1.0], [-1.0, 0.5, 0.25, 0.0], [0.5, 0.5, 0.5, 0.5], [1.5, -1.25, 0.0, 0.25]]
q = [0.5, -0.25, 0.75, 0.5]
for metric in ("l2", "ip"):
    v, i = topk(metric, q, xs, 2)
    cases.append(dict(name=f"quarters_{metric}", metric=metric, corpus=xs, query=q, k=2, scores=v, ids=i))

# LangChain relevance / threshold table: relevance = 1 - d/sqrt(2), kept when >= thr
s2 = math.sqrt(2)
rel = []
for d in (0.0, 0.5, 0.75 * s2, 1.0, 2.0):  # 0.75*sqrt(2) -> relevance exactly on the 0.25 boundary
    rel.append(dict(distance=d, euclidean=1.0 - d / s2, cosine=1.0 - d,
                    max_inner_product=(1.0 - d) if d > 0 else -1.0 * d))
relevance = dict(threshold=0.25, table=rel)

# normalise: zero row untouched, [3,4,0,0] -> [0.6,0.8,0,0]
normalize = dict(rows=[[3, 4, 0, 0], [0, 0, 0, 0], [0, 0, 0, 2]],
                 expect=[[0.6, 0.8, 0, 0], [0, 0, 0, 0], [0, 0, 0, 1]])

# remove_ids: positions 1 and 3 (duplicate 3, out-of-range 9) from 5 rows
remove = dict(rows=[[float(r)] * 4 for r in range(5)], ids=[3, 1, 3, 9], n_removed=2,
              expect=[[0.0] * 4, [2.0] * 4, [4.0] * 4])

with open(os.path.join(HERE, "kat_d4.json"), "w") as f:
    json.dump(dict(cases=cases, relevance=relevance, normalize=normalize, remove_ids=remove), f, indent=1)
print("wrote kat_d4.json:", len(cases), "search cases")
