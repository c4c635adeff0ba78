"""adaface_b200: B200-native (sm_100a) implementation of the AdaPrompt-conditioned U-Net transformer path of
zyt334/AdaFace - SpatialTransformer / BasicTransformerBlock / CrossAttention, AdaEmbedding + AttentionalPooler,
and the CFG combine + DDIM update - behind the reference's own module signatures.  See DESIGN.md."""
__version__ = "0.1.0"

from .attention import (BasicTransformerBlock, CrossAttention, FeedForward, GEGLU, Normalize,  # noqa: F401
                        SpatialTransformer)
from .ddim import DDIMSampler  # noqa: F401
from .embedding import AdaEmbedding, AttentionalPooler  # noqa: F401
