"""Golden vectors for the ingest / confusion-matrix rows, produced by the UNMODIFIED reference code:
``pad`` is taken verbatim (ast) from /root/reference/data_utils_SSL.py:47-54 (the module itself needs librosa / RawBoost to
import), ``calculate_confusion_matrix`` is imported from /root/reference/evaluate_metrics.py.  Test infrastructure only.

    python oracle/gen_ingest_golden.py  ->  tests/golden/ingest_kat.json
"""
import ast
import importlib.util
import json
import os
import sys

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
    # evaluate.py metadata loaders on a small synthetic protocol / metadata pair (duplicates, missing files, repeated metadata)
    import tempfile
    sys.path.insert(0, REF)   # evaluate.py imports its sibling evaluate_metrics
    spec = importlib.util.spec_from_file_location("ref_eval", os.path.join(REF, "evaluate.py"))
    ev = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ev)
    proto = ["DF_E_%07d" % i for i in (5, 3, 9, 3, 7, 1, 42)]
    meta = ["LA_00%02d DF_E_%07d mp3m4a asvspoof A%02d %s notrim eval vocoder - - - -" % (i, i, i, "spoof" if i % 3 else "bonafide")
            for i in (1, 2, 3, 5, 7, 9, 11)] + ["LA_0003 DF_E_0000003 mp3m4a asvspoof A03 spoof notrim eval vocoder - - - -"]
    with tempfile.TemporaryDirectory() as d:
        pp, mp = os.path.join(d, "proto.txt"), os.path.join(d, "meta.txt")
        open(pp, "w").write("\n".join(proto) + "\n")
        open(mp, "w").write("\n".join(meta) + "\n")
        kat = {"proto": proto, "meta": meta, "load_proto_file": ev.load_proto_file(pp), "load_metadata": ev.load_metadata(mp),
               "load_metadata_from_proto": ev.load_metadata_from_proto(mp, pp), "eval_dict": ev.eval_dict(mp)}
        # the evaluate.py command line itself (unmodified, as a subprocess): stdout on (a) the ragged pair above,
        # (b) a clean 400-trial pair with every label matched
        import subprocess
        rng = np.random.default_rng(7)

        def run_cli(tag, proto_l, meta_l, scores, thr):
            pf, mf, sf = (os.path.join(d, tag + e) for e in (".proto", ".meta", ".score"))
            open(pf, "w").write("\n".join(proto_l) + "\n")
            open(mf, "w").write("\n".join(meta_l) + "\n")
            lines = ["%s, %d" % (repr(float(v)), int(v > thr)) for v in scores]
            open(sf, "w").write("\n".join(lines) + "\n")
            out = subprocess.run([sys.executable, os.path.join(REF, "evaluate.py"), "--score_file", sf, "--protocol_file", pf,
                                  "--metadata_file", mf, "--threshold", repr(thr)], capture_output=True, text=True, check=True,
                                 cwd=REF).stdout
            return {"proto": proto_l, "meta": meta_l, "score_lines": lines, "threshold": thr, "stdout": out}

        n = 400
        names = ["LA_E_%07d" % i for i in rng.permutation(n)]
        lab = ["bonafide" if rng.random() < 0.2 else "spoof" for _ in range(n)]
        sc = [float(rng.normal(0.35 if l == "bonafide" else 0.6, 0.15)) for l in lab]
        sc[10] = sc[20]                                          # a tie across the two classes' neighbourhood
        meta_c = ["LA_%04d %s - - A%02d %s - eval" % (i % 50, nm, i % 19, l) for i, (nm, l) in enumerate(zip(names, lab))]
        kat["cli"] = {"ragged": run_cli("ragged", proto, meta, [0.1, 0.7, 0.4, 0.9, 0.55, 0.2, 0.65], 0.5),
                      "clean": run_cli("clean", names, meta_c, sc, 0.47)}
    json.dump(kat, open(os.path.join(ROOT, "tests/golden/evalproto_kat.json"), "w"), indent=1)
    print("wrote tests/golden/ingest_kat.json")


if __name__ == "__main__":
    main()
