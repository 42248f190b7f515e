"""occm_b200 -- B200-native (sm_100a) scoring hot path of nguyenvulong/occm.

waveform -> XLS-R / wav2vec 2.0 frontend -> AASIST backend -> one-class score -> EER,
behind the reference's Python surface (``models.sslassist.AModel``, ``losses.custom_loss``,
``evaluate_metrics.compute_eer``, the two score-file formats).  All compute lives in
``libocc_b200.so`` (C ABI: ``include/occ_b200.h``); see DESIGN.md and INTEGRATION.md.
"""
__version__ = "0.1.0"
