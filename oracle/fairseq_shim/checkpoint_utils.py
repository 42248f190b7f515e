import torch


class _Wav2Vec2Standin(torch.nn.Module):
    """fairseq Wav2Vec2Model's eval-time surface used by the reference: forward(x, mask=False, features_only=True) -> {'x': ...}."""

    def __init__(self):
        super().__init__()
        import torchaudio
        self.m = torchaudio.models.wav2vec2_xlsr_300m()

    def forward(self, x, mask=False, features_only=True):
        y, _ = self.m(x)
        return {"x": y}


def load_model_ensemble_and_task(paths):
    # the reference's hard-coded checkpoint path (models/sslassist.py:24) does not exist here: random-init weights, replaced by
    # the seeded synthetic checkpoint right after construction (oracle/ref_runner.py), as BASELINE.json's configs specify
    return [_Wav2Vec2Standin()], None, None
