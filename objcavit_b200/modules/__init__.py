"""Drop-in ``nn.Module`` surface of the head (same class names, constructor signatures, ``state_dict``
keys and return tuples as the reference's ``modules/`` package)."""
from .layers import PatchTransformerEncoder, PixelWiseDotProduct
from .miniViT import mViT
from .ObjCAViT import GridRandomPositionalEmbeddings, ObjCAViT, SelfAttnCrossAttn
from .heads import AdaBinsHead, GraphBinsHead, HostPipeline, build_head
from .decoder_handoff import FinalConvNHWC

__all__ = ["PatchTransformerEncoder", "PixelWiseDotProduct", "mViT", "GridRandomPositionalEmbeddings",
           "SelfAttnCrossAttn", "ObjCAViT", "AdaBinsHead", "GraphBinsHead", "HostPipeline", "build_head", "FinalConvNHWC"]
