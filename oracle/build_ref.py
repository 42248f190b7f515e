"""Recipe that makes the UNMODIFIED reference runnable away from /root/reference (test / benchmark infrastructure only).

    python -m oracle.build_ref            # in the build container, where /root/reference exists

The reference is 100 % Python, so there is nothing to compile: the recipe packs the files the scoring path needs
(`models/sslassist.py`, `evaluate_metrics.py`, `evaluate.py`, `calculate_eer.py`) byte-for-byte from where they lie under
/root/reference into ONE build artefact, `oracle/_ref/reference_py.zip` (git-ignored, NOT gpurun-ignored: it travels to the
GPU box like a built .so; the sources never enter this repository's tree or history), writes their SHA-256 into
`oracle/_ref/MANIFEST.json`, and puts the `fairseq` stand-in of SURVEY.md Appendix A beside it (`oracle/_ref/fairseq/`,
copied from `oracle/fairseq_shim/` -- our code, not the reference's: fairseq@a540213 cannot be installed here, torchaudio's
`wav2vec2_xlsr_300m()` is the in-image statement of the same network).
`oracle/ref_runner.py` then imports `models.sslassist` from the archive (zipimport) exactly as `oc_classifier.py:335` does.
`bench.py --impl reference` and `cpu_baseline` time that (`kind: "reference"`); without `oracle/_ref` they fall back to the
oracle port and say so (`kind: "port"`).
"""
from __future__ import annotations

import hashlib
import json
import os
import shutil
import sys
import zipfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"
DST = os.path.join(ROOT, "oracle", "_ref")
ARCHIVE = os.path.join(DST, "reference_py.zip")
FILES = ["models/sslassist.py", "models/xlsr.py", "models/lcnn.py", "models/senet.py", "models/occm.py",
         "evaluate_metrics.py", "evaluate.py", "calculate_eer.py"]


def build(verbose: bool = True) -> bool:
    """Returns True when oracle/_ref is (now) complete; False when /root/reference is absent and nothing was there before."""
    have = os.path.exists(ARCHIVE) and os.path.isdir(os.path.join(DST, "fairseq"))
    if not os.path.isdir(REF):
        return have
    os.makedirs(DST, exist_ok=True)
    manifest = {}
    with zipfile.ZipFile(ARCHIVE, "w", zipfile.ZIP_DEFLATED) as z:
        for f in FILES:
            data = open(os.path.join(REF, f), "rb").read()
            z.writestr(zipfile.ZipInfo(f, date_time=(2024, 1, 1, 0, 0, 0)), data)
            manifest[f] = hashlib.sha256(data).hexdigest()
        z.writestr(zipfile.ZipInfo("models/__init__.py", date_time=(2024, 1, 1, 0, 0, 0)), b"")
    shim_src = os.path.join(ROOT, "oracle", "fairseq_shim")
    shim_dst = os.path.join(DST, "fairseq")
    if os.path.isdir(shim_dst):
        shutil.rmtree(shim_dst)
    shutil.copytree(shim_src, shim_dst)
    sub = os.path.join(REF, ".SUBMODULES.json")
    manifest["_reference_commit"] = json.load(open(sub)) if os.path.exists(sub) else None
    with open(os.path.join(DST, "MANIFEST.json"), "w") as fh:
        json.dump(manifest, fh, indent=1)
    if verbose:
        print("oracle/_ref built:", ", ".join(FILES))
    return True


if __name__ == "__main__":
    ok = build()
    sys.exit(0 if ok else 1)
