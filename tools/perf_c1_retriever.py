"""Developer tool (GPU box): the C1 retriever call (10k chunks, k=4, threshold 0.25) split by kind of query --
planted next to a chunk (one hit kept) and unrelated (nothing kept) -- with its parts timed separately."""
import os, sys, json, time, warnings
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from generativeaiexamples_b200 import create_vectorstore_langchain, synth

d, n = 1024, 10_000
xs = synth.corpus_host(n, d)
q, _ = synth.queries_host(64, d, xs, noise=0.6)


class Emb:
    def __init__(self):
        self.table = {}

    def embed_documents(self, texts):
        return [self.table[t] for t in texts]

    def embed_query(self, t):
        return self.table[t]


emb = Emb()
texts = [f"chunk {i}" for i in range(n)]
for t_, v in zip(texts, xs):
    emb.table[t_] = v.tolist()
qt = [f"query {i}" for i in range(64)]
for t_, v in zip(qt, q):
    emb.table[t_] = v.tolist()
vs = create_vectorstore_langchain(emb, name="b200flat", dimensions=d)
vs.add_texts(texts, [{"source": f"f{i % 20}.pdf"} for i in range(n)])
retr = vs.as_retriever(search_type="similarity_score_threshold", search_kwargs={"score_threshold": 0.25, "k": 4})


def p50(fn, idx, reps=400):
    lat = []
    for i in range(reps):
        a = idx[i % len(idx)]
        t0 = time.perf_counter()
        fn(a)
        lat.append((time.perf_counter() - t0) * 1e6)
    lat = sorted(lat[50:])
    return round(lat[len(lat) // 2], 1)


rec = {}
with warnings.catch_warnings():
    warnings.simplefilter("ignore")
    planted, unrelated = list(range(0, 64, 2)), list(range(1, 64, 2))
    rec["retriever_planted_us"] = p50(lambda i: retr.get_relevant_documents(qt[i]), planted)
    rec["retriever_unrelated_us"] = p50(lambda i: retr.get_relevant_documents(qt[i]), unrelated)
    rec["with_relevance_scores_us"] = p50(lambda i: vs.similarity_search_with_relevance_scores(qt[i], k=4, score_threshold=0.25), planted)
    rec["engine_threshold_call_us"] = p50(lambda i: vs.index.search_threshold(q[i: i + 1], 4, "EUCLIDEAN_DISTANCE", 0.25), planted)
    rec["engine_search_call_us"] = p50(lambda i: vs.index.search(q[i: i + 1], 4), planted)
    vec = [emb.table[t_] for t_ in qt]
    rec["by_vector_us"] = p50(lambda i: vs.similarity_search_with_score_by_vector(vec[i], k=4), planted)
print(json.dumps(rec))
