"""``models/xlsr.py`` of the reference (``/root/reference/models/xlsr.py:25-52``): the stand-alone
XLS-R feature extractor.  Same class as ``models.sslassist.SSLModel``; ``forward`` = ``extract_feat``."""
from .sslassist import SSLModel  # noqa: F401
