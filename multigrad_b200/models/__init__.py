"""Model families: the native fused soft-binned models and the reference's
example model written against the generic API."""
from .erf_binned import ErfBinnedModel, SMFModel, AdamFit  # noqa: F401
from .linear_link import LinearLinkBinnedModel  # noqa: F401
from .linked import LinkedBinnedModel  # noqa: F401
from . import smf  # noqa: F401
from .smf import MySMFModel, TutorialSMFModel, ParamTuple  # noqa: F401
