"""Model surface of the reference's ``models/`` package that lies on the scoring hot path."""
from .sslassist import AModel, Model, SSLModel  # noqa: F401
