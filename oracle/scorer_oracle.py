"""CPU oracle for the CassMantle guess scorer -- TEST INFRASTRUCTURE, NOT PRODUCT.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline /
``--impl reference`` legs may import this module.  Nothing under
``cassmantle_b200/`` imports it; the product path is CUDA only and raises when the
extension is missing.

PARITY UNPINNED.  The reference has no tests, fixtures or golden vectors
 and its scorer cannot be imported here (gensim, aioredis,
nltk are not installed; the 3.6 GB weights file is absent).  The arithmetic lives
in a third-party dependency that is NOT under /root/reference:

    gensim  (requirements.txt:5, unpinned, no lock file; API of the 4.x line)

so this file restates (a) the reference's own control flow,
``src/backend.py:303-317``, line by line, and (b) gensim's published algorithm
for ``KeyedVectors.similarity`` / ``matutils.unitvec``:

    similarity(w1, w2) = dot(unitvec(vectors[key_to_index[w1]]),
                             unitvec(vectors[key_to_index[w2]]))
    unitvec(v)         = v                       if nrm2(v) == 0
                         scal(1/nrm2(v), v).astype(v.dtype)   otherwise
    where nrm2 / scal are scipy's *double precision* BLAS routines
    (``blas('nrm2', np.array([], dtype=float))``), so an fp32 row is promoted,
    its norm and the scaling are done in fp64, and the result is rounded back
    to fp32 before an fp32 ``numpy.dot``.

scipy is installed here, so ``unitvec`` below calls the very same BLAS entry
points gensim binds.  The outputs are pinned only by the known-answer cases in
``tests/golden/scorer_kat.json`` (authored in this repo from the semantics
above; generator: ``tests/golden/make_golden.py``).  What IS checked against the reference's
own code, live, is part (a): ``tests/test_reference_backend.py`` imports ``src/backend.py``
unmodified and runs its ``compute_score`` / ``compute_scores`` with ``OracleWordVectors`` as
``self.wv``; only the arithmetic inside gensim stays a restatement.

Extension (labelled, SURVEY.md §0 Track A): a guess of N > 1 words is mean-pooled
exactly like gensim's classic ``n_similarity`` -- ``array(rows).mean(axis=0)`` ->
``unitvec`` -> ``dot`` -- which the reference itself never calls (its frontend
forbids multi-word input, static/script.js:418).  N = 1 reduces to ``similarity``
bit for bit.  Any out-of-vocabulary word raises KeyError -> ``min_score``, the same
way ``self[key]`` would.
"""
from __future__ import annotations

import asyncio
from decimal import ROUND_HALF_UP, Decimal
from typing import Dict, List, Optional, Sequence

import numpy as np
from scipy.linalg import get_blas_funcs

# gensim/matutils.py: blas_nrm2 = blas('nrm2', np.array([], dtype=float)); same for scal.
_blas_nrm2 = get_blas_funcs(("nrm2",), (np.array([], dtype=float),))[0]
_blas_scal = get_blas_funcs(("scal",), (np.array([], dtype=float),))[0]


def unitvec(vec: np.ndarray) -> np.ndarray:
    """gensim ``matutils.unitvec`` for a dense array, norm='l2'."""
    if vec.size == 0:
        return vec
    veclen = _blas_nrm2(vec)
    if veclen > 0.0:
        return _blas_scal(1.0 / veclen, vec).astype(vec.dtype)
    return vec


def similarity(table: np.ndarray, i: int, j: int) -> np.float32:
    """gensim ``KeyedVectors.similarity`` on row indices (call site src/backend.py:307)."""
    return np.dot(unitvec(table[i]), unitvec(table[j]))


def n_similarity(table: np.ndarray, rows1: Sequence[int], rows2: Sequence[int]) -> np.float32:
    """gensim classic ``KeyedVectors.n_similarity``: mean of rows -> unitvec -> dot."""
    if not (len(rows1) and len(rows2)):
        raise ZeroDivisionError("At least one of the passed list is empty.")
    v1 = [table[r] for r in rows1]
    v2 = [table[r] for r in rows2]
    return np.dot(unitvec(np.array(v1).mean(axis=0)), unitvec(np.array(v2).mean(axis=0)))


class OracleWordVectors:
    """Minimal stand-in for the ``KeyedVectors`` object at src/backend.py:45."""

    def __init__(self, vectors: np.ndarray, keys: Sequence[str]) -> None:
        assert vectors.dtype == np.float32 and vectors.ndim == 2
        self.vectors = vectors
        self.key_to_index: Dict[str, int] = {}
        for i, k in enumerate(keys):          # first occurrence wins, like gensim's add_vectors
            self.key_to_index.setdefault(k, i)

    def __getitem__(self, key: str) -> np.ndarray:
        return self.vectors[self.key_to_index[key]]   # KeyError when absent

    def similarity(self, w1: str, w2: str) -> np.float32:
        return np.dot(unitvec(self[w1]), unitvec(self[w2]))

    def n_similarity(self, ws1: Sequence[str], ws2: Sequence[str]) -> np.float32:
        if not (len(ws1) and len(ws2)):
            raise ZeroDivisionError("At least one of the passed list is empty.")
        v1 = [self[k] for k in ws1]
        v2 = [self[k] for k in ws2]
        return np.dot(unitvec(np.array(v1).mean(axis=0)), unitvec(np.array(v2).mean(axis=0)))


class OracleBackend:
    """The scorer half of ``Backend`` (src/backend.py:303-317), nothing else."""

    def __init__(self, wv: OracleWordVectors, min_score: float = 0.01) -> None:
        self.wv = wv
        self.min_score = min_score            # live value 0.01, src/server.py:17

    def compute_score(self, inputs: str, answer: str):
        inputs, answer = inputs.lower(), answer.lower()          # backend.py:304
        if inputs == answer:                                      # backend.py:305
            return 1.0
        try:
            score = self.wv.similarity(inputs, answer)            # backend.py:307
        except KeyError:
            score = self.min_score                                # backend.py:309
        return max(self.min_score, score)                         # backend.py:310

    async def compute_scores(self, data: Dict[str, Dict[str, str]]) -> Dict[str, str]:
        scores: Dict[str, str] = {}
        for key in data.keys():                                   # backend.py:314
            score = self.compute_score(data[key]["input"], data[key]["answer"])
            scores.update({key: str(score)})                      # backend.py:316
        return scores

    def compute_scores_sync(self, data: Dict[str, Dict[str, str]]) -> Dict[str, str]:
        return asyncio.run(self.compute_scores(data))

    # ---- labelled extension: a guess that is a phrase of several words -------------
    def compute_score_phrase(self, inputs: str, answer: str):
        inputs, answer = inputs.lower(), answer.lower()
        if inputs == answer:
            return 1.0
        try:
            score = self.wv.n_similarity(inputs.split(), [answer])
        except KeyError:
            score = self.min_score
        return max(self.min_score, score)


# --------------------------------------------------------------------------- #
# Index-form batch scorers: same arithmetic, inputs as the CUDA kernel sees them.
# --------------------------------------------------------------------------- #
def score_batch_faithful(table: np.ndarray, guess_idx: np.ndarray, guess_len: np.ndarray,
                         answer_rows: np.ndarray, answer_slot: np.ndarray,
                         exact: Optional[np.ndarray], min_score: float) -> np.ndarray:
    """Per-pair Python loop, one ``similarity`` call per pair -- how the reference runs.

    Returns fp32[B].  ``max(min_score, score)`` follows Python semantics: the cosine is
    kept only when ``score > min_score`` (NaN therefore yields ``min_score``)."""
    B = guess_idx.shape[0]
    out = np.empty(B, dtype=np.float32)
    for b in range(B):
        if exact is not None and exact[b]:
            out[b] = 1.0
            continue
        n = int(guess_len[b])
        rows = guess_idx[b, :n]
        a = int(answer_rows[int(answer_slot[b])])
        try:
            if n <= 0:
                raise KeyError("empty guess")      # '' is never a key of the table
            if (rows < 0).any() or a < 0:
                raise KeyError("out of vocabulary")
            if n == 1:
                score = similarity(table, int(rows[0]), a)
            else:
                score = n_similarity(table, rows, [a])
        except KeyError:
            score = min_score
        out[b] = max(min_score, score)
    return out


def score_batch_vectorised(table: np.ndarray, guess_idx: np.ndarray, guess_len: np.ndarray,
                           answer_rows: np.ndarray, answer_slot: np.ndarray,
                           exact: Optional[np.ndarray], min_score: float) -> np.ndarray:
    """The "strong" CPU baseline: same math, batched numpy, BLAS threads."""
    B, N = guess_idx.shape
    valid = np.arange(N)[None, :] < guess_len[:, None]
    oov = ((guess_idx < 0) & valid).any(axis=1) | (guess_len <= 0)
    safe = np.where(valid & (guess_idx >= 0), guess_idx, 0)
    acc = np.zeros((B, table.shape[1]), dtype=np.float32)
    for k in range(N):                                   # in-order fp32 sum == mean(axis=0)
        acc += table[safe[:, k]] * valid[:, k, None].astype(np.float32)
    mean = acc / np.maximum(guess_len, 1).astype(np.float32)[:, None]
    gn = np.sqrt(np.einsum("bd,bd->b", mean.astype(np.float64), mean.astype(np.float64)))
    gu = np.where(gn[:, None] > 0, mean / np.where(gn > 0, gn, 1.0)[:, None], mean).astype(np.float32)
    ans = table[np.maximum(answer_rows, 0)]
    an = np.sqrt(np.einsum("ad,ad->a", ans.astype(np.float64), ans.astype(np.float64)))
    au = np.where(an[:, None] > 0, ans / np.where(an > 0, an, 1.0)[:, None], ans).astype(np.float32)
    cos = np.einsum("bd,bd->b", gu, au[answer_slot]).astype(np.float32)
    oov |= answer_rows[answer_slot] < 0
    m32 = np.float32(min_score)
    out = np.where(~oov & (cos > m32), cos, m32).astype(np.float32)
    if exact is not None:
        out = np.where(exact.astype(bool), np.float32(1.0), out)
    return out


# --------------------------------------------------------------------------- #
# Output formatting and what the consumers do with it.
# --------------------------------------------------------------------------- #
def format_score(score) -> str:
    """``str(score)`` exactly as src/backend.py:316 produces it (Python float or np.float32)."""
    return str(score)


def js_to_fixed2_of_percent(score_str: str) -> str:
    """static/script.js:297-303: ``(parseFloat(s) * 100).toFixed(2)`` emulated exactly.

    parseFloat gives the nearest double; ``* 100`` is one fp64 multiply; toFixed(2) picks
    the integer n minimising |n/100 - x| on the exact binary value, larger n on ties."""
    x = float(score_str) * 100.0
    q = Decimal(x).quantize(Decimal("0.01"), rounding=ROUND_HALF_UP)
    return f"{q:.2f}"


def centi_score(score_str: str) -> int:
    """Displayed percentage in hundredths: ``round(score * 1e4)`` the way the browser rounds."""
    return int(Decimal(js_to_fixed2_of_percent(score_str)) * 100)


def placeholder(score_str: str) -> str:
    """The text the browser shows for one mask (static/script.js:297-308)."""
    s = float(score_str)
    if s > 0:
        return js_to_fixed2_of_percent(score_str) if s > 0.1 else "try again"
    return ""


def score_to_blur(score: float, min_blur: float = 0.0, max_blur: float = 15) -> float:
    """src/backend.py:319-320."""
    return min_blur + (1 - score ** 2) * (max_blur - min_blur)


def client_result(scores: Dict[str, str]) -> Dict[str, object]:
    """What src/server.py:78-89 derives from the scorer's strings (mean, won)."""
    mean_score = sum([float(s) for s in scores.values()]) / len(scores)
    return {"mean": mean_score, "won": int(mean_score == 1)}


def pairs_to_batch(key_to_index: Dict[str, int], data: Dict[str, Dict[str, str]]):
    """Host-side lowering of a compute_scores payload to index form (test helper)."""
    keys = list(data.keys())
    g = np.full((len(keys), 1), -1, dtype=np.int32)
    a = np.full(len(keys), -1, dtype=np.int32)
    ex = np.zeros(len(keys), dtype=np.uint8)
    for n, k in enumerate(keys):
        i, w = data[k]["input"].lower(), data[k]["answer"].lower()
        ex[n] = i == w
        g[n, 0] = key_to_index.get(i, -1)
        a[n] = key_to_index.get(w, -1)
    return keys, g, np.ones(len(keys), dtype=np.int32), a, np.arange(len(keys), dtype=np.int32), ex


# --------------------------------------------------------------------------- #
# Vocabulary lookup: the reference's mechanism IS a Python dict.
# --------------------------------------------------------------------------- #
def lookup_rows(index_to_key: Sequence[str], words: Sequence[str]) -> np.ndarray:
    """``key_to_index.get(word.lower(), -1)`` with gensim's first-occurrence-wins dict
    (src/backend.py:304 + the lookup inside wv.similarity at src/backend.py:307)."""
    k2i: Dict[str, int] = {}
    for i, k in enumerate(index_to_key):
        k2i.setdefault(k, i)
    return np.asarray([k2i.get(w.lower(), -1) for w in words], dtype=np.int32)


# --------------------------------------------------------------------------- #
# most_similar: src/backend.py:297-301 -> gensim KeyedVectors.most_similar
# --------------------------------------------------------------------------- #
def most_similar_dists(table: np.ndarray, row: int) -> np.ndarray:
    """gensim 4.x, positive=[word]: dists = dot(vectors, mean) / norms with
    norms = np.linalg.norm(vectors, axis=1) and mean = unitvec(vectors[row] / norms[row])."""
    norms = np.linalg.norm(table, axis=1)
    with np.errstate(divide="ignore", invalid="ignore"):
        mean = unitvec(np.array([table[row] / norms[row]]).mean(axis=0)).astype(np.float32)
        return np.dot(table, mean) / norms


def most_similar_rows(table: np.ndarray, row: int, topn: int = 50):
    """[(row, float(sim))] best first, the query row dropped (gensim asks for topn + 1)."""
    dists = most_similar_dists(table, row)
    x = -dists                                              # matutils.argsort(..., reverse=True)
    k = topn + 1
    if k >= x.size:
        best = np.argsort(x)[:k]
    else:
        part = np.argpartition(x, k)[:k]
        best = part.take(np.argsort(x.take(part)))
    return [(int(r), float(dists[r])) for r in best if int(r) != row][:topn]
