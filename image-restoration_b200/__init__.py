"""b200ir — B200-native (sm_100a) Restormer-family transformer block behind the reference's nn.Module API."""
from . import _lib, ops  # noqa: F401  (fails loudly if libb200ir.so is missing)
from .modules import (Attention, ElementwiseGatedAttention, GatedSpatialFFN, BiasFree_LayerNorm, CrossAttention, GatedMDTA, GatedTransformerBlock, Downsample, EncoderBlock, EncoderResidualGroup, FeedForward, LayerNorm, OverlapPatchEmbed,  # noqa: F401
                      Restormer, TransformerBlock, Upsample, WithBias_LayerNorm)

from .dropin import accelerate  # noqa: F401
from . import pipeline  # noqa: F401
from .pipeline import crop_augment_degrade, psnr_ssim, tiled_restore  # noqa: F401
from .adair import AdaIR, Chanel_Cross_Attention, ChannelGate, FreModule, FreRefine, SpatialGate  # noqa: F401
from .moce import (AdapterLayer, DecoderBlock, DecoderResidualGroup, FFTAttention, FrequencyEmbedding, HighPassConv2d,  # noqa: F401
                   ModExpert, MoCEIR, RoutingFunction)
from .autograd_ops import invalidate_caches  # noqa: F401
from .training import AdamW, GraphedForward, GraphedTrainStep, L1Loss, PipelinedForward, allreduce_flat_grads, train_step  # noqa: F401

__all__ = ["AdaIR", "accelerate", "pipeline", "crop_augment_degrade", "psnr_ssim", "tiled_restore", "FreModule", "FreRefine", "SpatialGate", "ChannelGate", "Chanel_Cross_Attention", "ElementwiseGatedAttention", "GatedSpatialFFN", "MoCEIR", "DecoderBlock", "DecoderResidualGroup", "AdapterLayer", "ModExpert", "FFTAttention", "RoutingFunction", "FrequencyEmbedding", "HighPassConv2d", "invalidate_caches", "AdamW", "GraphedForward", "PipelinedForward", "GraphedTrainStep", "L1Loss", "allreduce_flat_grads", "train_step", "Attention", "BiasFree_LayerNorm", "CrossAttention", "GatedMDTA", "GatedTransformerBlock", "Downsample", "EncoderBlock", "EncoderResidualGroup", "FeedForward", "LayerNorm", "OverlapPatchEmbed",
           "Restormer", "TransformerBlock", "Upsample", "WithBias_LayerNorm", "ops"]
