"""Stand-in for the `fairseq` package the reference imports (`models/sslassist.py:9,25`), SURVEY.md Appendix A.

fairseq@a54021305d6b (requirement.txt:33) is not installable in this image (no network).  The reference only calls
`fairseq.checkpoint_utils.load_model_ensemble_and_task([path])` and then
`model(x, mask=False, features_only=True)['x']`; torchaudio's `wav2vec2_xlsr_300m()` is the same XLS-R-300M network
(torchaudio documents numerical equivalence with fairseq and ships the key map, torchaudio/utils/import_fairseq.py).
Test / benchmark infrastructure only: copied to `oracle/_ref/fairseq/` by `oracle/build_ref.py`.
"""
from . import checkpoint_utils  # noqa: F401
