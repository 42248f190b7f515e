"""Golden vectors for the ingest / confusion-matrix rows, produced by the UNMODIFIED reference code:
``pad`` is taken verbatim (ast) from /root/reference/data_utils_SSL.py:47-54 (the module itself needs librosa / RawBoost to
import), ``calculate_confusion_matrix`` is imported from /root/reference/evaluate_metrics.py.  Test infrastructure only.

    python oracle/gen_ingest_golden.py  ->  tests/golden/ingest_kat.json
"""
import ast
import importlib.util
import json
import os

import numpy as np

REF = "/root/reference"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def reference_pad():
    src = open(os.path.join(REF, "data_utils_SSL.py")).read()
    tree = ast.parse(src)
    fn = next(n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name == "pad")
    ns = {"np": np}
    exec(compile(ast.Module(body=[fn], type_ignores=[]), "data_utils_SSL.py", "exec"), ns)
    return ns["pad"]


def main():
    pad = reference_pad()
    spec = importlib.util.spec_from_file_location("ref_em", os.path.join(REF, "evaluate_metrics.py"))
    em = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(em)
    rng = np.random.default_rng(7)
    cases = []
    for n, max_len in ((5, 12), (12, 12), (20, 12), (1, 7), (7, 64), (64600, 64600), (3000, 64600)):
        x = rng.integers(-32768, 32768, n, dtype=np.int16)
        xf = (x.astype(np.float32) / np.float32(32768.0)).astype(np.float32)
        y = pad(xf, max_len)
        item = {"n": n, "max_len": max_len, "seed_note": "rng(7) sequential", "sha_in": None}
        if n <= 64:
            item["x"] = x.tolist()
            item["y"] = [float(v) for v in y]
        else:
            item["x_first"] = x[:8].tolist()
            item["y_sum"] = float(np.sum(y.astype(np.float64)))
            item["y_tail"] = [float(v) for v in y[-4:]]
        cases.append(item)
    t = rng.normal(1, 1, 1000).astype(np.float32).astype(np.float64)
    nt = rng.normal(-1, 1, 3000).astype(np.float32).astype(np.float64)
    cm = []
    for thr in (-10.0, 0.0, 0.0025480177719146013, float(t[3]), 10.0):
        cm.append({"thr": thr, "counts": [int(v) for v in em.calculate_confusion_matrix(t, nt, thr)]})
    out = {"pad_cases": cases, "cm_seed": "rng(7) after the pad cases: normal(1,1,1000), normal(-1,1,3000), fp32->fp64", "cm": cm}
    json.dump(out, open(os.path.join(ROOT, "tests/golden/ingest_kat.json"), "w"), indent=1)
    print("wrote tests/golden/ingest_kat.json")


if __name__ == "__main__":
    main()
