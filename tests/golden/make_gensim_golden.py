"""Pins the scorer's arithmetic to the REAL gensim, on any machine that has gensim.

    python tests/golden/make_gensim_golden.py --reference /path/to/CassMantle [--out FILE]

The reference's scorer (src/backend.py:297-317) does its arithmetic inside gensim
(requirements.txt:5), which cannot be installed on the build or GPU boxes of this repo, so
oracle/scorer_oracle.py is a restatement of gensim's published algorithm and its header says
"parity unpinned".  This script closes that gap wherever gensim exists -- the 3.6 GB
word2vec-google-news-300 download is NOT needed, because what has to be pinned is arithmetic,
not weights:

  1. builds a real ``gensim.models.KeyedVectors`` from this repo's seeded synthetic table
     (cassmantle_b200/synth.py: table_rows + make_vocab) plus the 12 hand-built known-answer rows
     of tests/golden/scorer_kat.json;
  2. imports the reference's ``src/backend.py`` UNMODIFIED from --reference (aioredis / nltk are
     replaced by empty stand-ins when absent: neither is executed by the methods used), sets
     ``self.wv`` to that KeyedVectors and ``self.min_score = 0.01`` (src/server.py:17);
  3. runs the reference's own ``compute_score`` on a few thousand (guess, answer) string pairs
     (the known-answer cases, case folding, unknown words, near and far rows),
     ``compute_scores`` on request-shaped dicts and ``most_similar`` on a few words;
  4. writes tests/golden/scorer_gensim_golden.json: every result as a hex float plus its Python
     type and ``str()``, with the gensim / numpy / scipy versions and numpy's BLAS recorded.

tests/test_gensim_golden.py consumes that file when it is present (CPU: the oracle against it;
``-m gpu``: the CUDA path against it) and skips with a clear reason when it is not.  Commit the
JSON after running this script and the oracle's header can say "parity pinned".

``--dry-run`` replaces step 1's KeyedVectors by the oracle's own OracleWordVectors, so that the
whole kit (generation, file format, consumers) can be exercised without gensim; a dry-run file
is marked as such and pins nothing.
"""
from __future__ import annotations

import argparse
import asyncio
import json
import os
import sys
import types

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from cassmantle_b200 import synth            # noqa: E402

SYNTH_ROWS, SYNTH_SEED = 20_000, 0
MIN_SCORE = 0.01
DEFAULT_OUT = os.path.join(ROOT, "tests", "golden", "scorer_gensim_golden.json")
KAT_PATH = os.path.join(ROOT, "tests", "golden", "scorer_kat.json")


def kat_rows():
    """The hand-built known-answer rows (same decoding as tests/util.py: dense_kat_table)."""
    with open(KAT_PATH) as f:
        g = json.load(f)
    rows = []
    for k in g["kat"]["keys"]:
        desc = g["kat"]["rows"][k]
        v = np.zeros(g["dim"], dtype=np.float32)
        if isinstance(desc, list):
            v[: desc[2]] = np.float32(desc[1])
        else:
            for i, x in desc.items():
                v[int(i)] = np.float32(x)
        rows.append(v)
    return g, ["kat_" + k for k in g["kat"]["keys"]], np.stack(rows)


def build_table():
    """(keys, fp32 matrix): the synthetic rows under their make_vocab keys, then the KAT rows
    under 'kat_<name>'.  Deterministic: a consumer rebuilds exactly this."""
    keys = synth.make_vocab(SYNTH_ROWS, seed=SYNTH_SEED)
    table = synth.table_rows(np.arange(SYNTH_ROWS), seed=SYNTH_SEED)
    _, kat_keys, kat = kat_rows()
    assert not set(kat_keys) & set(keys)
    return keys + kat_keys, np.concatenate([table, kat])


def build_pairs(keys):
    """A few thousand (guess, answer) strings: every class of src/backend.py:303-310."""
    g, kat_keys, _ = kat_rows()
    pairs = [("kat_" + c["input"] if c["input"].lower() in g["kat"]["keys"] else c["input"],
              "kat_" + c["answer"] if c["answer"].lower() in g["kat"]["keys"] else c["answer"])
             for c in g["kat"]["cases"]]
    pairs += [(a, b) for a in kat_keys for b in kat_keys]                     # all KAT rows against each other
    rng = np.random.default_rng(20261107)
    n = SYNTH_ROWS
    clu = synth.cluster_of(np.arange(n), SYNTH_SEED)
    answers = [int(x) for x in rng.choice(n, size=24, replace=False)]
    for a in answers:
        same = np.nonzero(clu == clu[a])[0]
        for r in rng.choice(same, size=min(40, same.size), replace=False):     # near: cosine ~0.5
            pairs.append((keys[int(r)], keys[a]))
        for r in rng.integers(0, n, size=100):                                 # far: cosine ~0, many floored
            pairs.append((keys[int(r)], keys[a]))
        pairs.append((keys[a].upper(), keys[a]))                              # equal after lower() -> 1.0
        pairs.append((keys[a].capitalize(), keys[int(rng.integers(0, n))]))    # case folding, then a lookup
        pairs.append(("no-such-word", keys[a]))                               # KeyError -> min_score
        pairs.append((keys[a], "No Such Answer"))
    pairs += [("", ""), ("", keys[0]), (" ", keys[0]), ("kat_void", "kat_void"), ("kat_void", keys[3])]
    return pairs


def import_reference_backend(reference_root: str):
    """The reference's src/backend.py, unmodified.  aioredis / nltk / gensim are replaced by empty
    stand-ins only when they are not importable (none of them is executed by the scorer methods;
    the gensim OBJECT used is the real one, built by the caller)."""
    if not os.path.isfile(os.path.join(reference_root, "src", "backend.py")):
        raise SystemExit(f"--reference {reference_root!r} has no src/backend.py (a checkout of SnowCheetos/CassMantle)")
    stand_ins = []
    for name in ("aioredis", "aioredis.exceptions", "nltk", "gensim", "gensim.models"):
        try:
            __import__(name)
        except Exception:       # noqa: BLE001
            sys.modules[name] = types.ModuleType(name)
            stand_ins.append(name)
    if "aioredis.exceptions" in stand_ins:
        sys.modules["aioredis.exceptions"].LockError = type("LockError", (Exception,), {})
    if "aioredis" in stand_ins:
        sys.modules["aioredis"].Redis = object
    if "gensim.models" in stand_ins:
        sys.modules["gensim.models"].KeyedVectors = object
    for m in [k for k in sys.modules if k == "src" or k.startswith("src.")]:
        del sys.modules[m]
    sys.path.insert(0, reference_root)
    try:
        from src.backend import Backend
    finally:
        sys.path.remove(reference_root)
        # leave the interpreter as it was found (this function is also called from the tests)
        for m in stand_ins + [k for k in sys.modules if k == "src" or k.startswith("src.")]:
            sys.modules.pop(m, None)
    return Backend, stand_ins


def _enc(x):
    """A scorer result, losslessly: Python type, str() (what compute_scores returns), hex value."""
    return {"type": type(x).__name__, "str": str(x), "hex": float(x).hex()}


def generate(wv, reference_root: str, out_path: str, versions: dict, dry_run: bool) -> dict:
    Backend, stand_ins = import_reference_backend(reference_root)
    ref = object.__new__(Backend)           # __init__ wants api_key.txt and the 3.6 GB file
    ref.min_score = MIN_SCORE               # src/server.py:17
    ref.wv = wv
    keys, _ = build_table()
    pairs = build_pairs(keys)
    scores = [_enc(ref.compute_score(a, b)) for a, b in pairs]                 # src/backend.py:303-310
    requests = []
    for i in range(0, 400, 2):                                                  # src/backend.py:312-317
        data = {str(3 + i): {"input": pairs[i][0], "answer": pairs[i][1]},
                str(11 + i): {"input": pairs[i + 1][0], "answer": pairs[i + 1][1]}}
        requests.append({"data": data, "result": asyncio.run(ref.compute_scores(data))})
    similar = []
    for w, topn in ((keys[5], 50), (keys[1234].upper(), 10), (keys[19_999], 200), ("kat_ones", 5)):
        got = ref.most_similar(w, topn=topn)                                    # src/backend.py:297-301
        similar.append({"word": w, "topn": topn, "result": [[k, float(s).hex()] for k, s in got]})
    try:
        ref.most_similar("No-Such-Word")
        unknown = None
    except Exception as exc:       # noqa: BLE001
        unknown = {"type": type(exc).__name__, "arg0": exc.args[0]}
    doc = {
        "what": "outputs of the reference's own Backend.compute_score / compute_scores / most_similar "
                "(src/backend.py:297-317) with self.wv = a gensim KeyedVectors over this repo's synthetic table",
        "dry_run": bool(dry_run),
        "versions": versions, "stand_ins": stand_ins, "min_score": MIN_SCORE,
        "table": {"synthetic_rows": SYNTH_ROWS, "seed": SYNTH_SEED, "kat_rows_appended": True,
                  "n_keys": len(keys)},
        "pairs": [list(p) for p in pairs], "scores": scores, "requests": requests,
        "most_similar": similar, "most_similar_unknown_word": unknown,
    }
    with open(out_path, "w") as f:
        json.dump(doc, f, indent=0)
        f.write("\n")
    return doc


def real_keyed_vectors():
    import gensim
    import scipy
    from gensim.models import KeyedVectors
    keys, table = build_table()
    kv = KeyedVectors(vector_size=table.shape[1])
    kv.add_vectors(keys, table)
    blas = None
    try:
        blas = str(np.show_config(mode="dicts")["Build Dependencies"]["blas"])
    except Exception:       # noqa: BLE001
        pass
    versions = {"gensim": gensim.__version__, "numpy": np.__version__, "scipy": scipy.__version__,
                "python": sys.version.split()[0], "numpy_blas": blas}
    return kv, versions


def oracle_word_vectors():
    """--dry-run: the oracle's stand-in for KeyedVectors, with gensim's most_similar return type."""
    import scipy
    from oracle import scorer_oracle as O

    class WordVectors(O.OracleWordVectors):
        def most_similar(self, word, topn=10):
            names = list(self.key_to_index)
            return [(names[r], s) for r, s in O.most_similar_rows(self.vectors, self.key_to_index[word], topn)]

    keys, table = build_table()
    versions = {"gensim": None, "numpy": np.__version__, "scipy": scipy.__version__,
                "python": sys.version.split()[0], "numpy_blas": None}
    return WordVectors(table, keys), versions


def main() -> None:
    ap = argparse.ArgumentParser(description=__doc__.split("\n\n")[0])
    ap.add_argument("--reference", default="/root/reference", help="checkout of SnowCheetos/CassMantle")
    ap.add_argument("--out", default=DEFAULT_OUT)
    ap.add_argument("--dry-run", action="store_true", help="use the oracle instead of gensim (pins nothing)")
    args = ap.parse_args()
    if args.dry_run:
        wv, versions = oracle_word_vectors()
    else:
        try:
            wv, versions = real_keyed_vectors()
        except ImportError as exc:
            raise SystemExit(f"gensim is not importable here ({exc}); run this on a machine that has it, "
                             f"or pass --dry-run to exercise the kit without pinning anything")
    doc = generate(wv, args.reference, args.out, versions, args.dry_run)
    print(f"wrote {args.out}: {len(doc['pairs'])} pairs, {len(doc['requests'])} requests, "
          f"{len(doc['most_similar'])} most_similar queries, gensim {versions['gensim']}")


if __name__ == "__main__":
    main()
