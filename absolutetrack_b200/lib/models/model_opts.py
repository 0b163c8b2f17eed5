"""Architecture options (mirror of the reference's lib/models/model_opts.py:11-39).

The CUDA library hard-wires exactly this configuration (csrc/model.cu layer table); the
dataclass exists so callers that construct or inspect ModelOpts keep working."""
from dataclasses import dataclass


@dataclass
class ModelOpts:
    network: str = "resnet_layers_2352-f32"
    nImageFeatureChannels: int = 72
    nSkeletonFeatureChannels: int = 4
    nTemporalMemoryChannels: int = 18
    useUnscaledAsCanonical: bool = False
    nMultiViewFusionBlocks: int = 2
    nTemporalBlocks: int = 3
    nPoseRegressionBlocks: int = 2
    spatialFTLRatio: float = 1.0
    temporalFTLRatio: float = 1.0
    nWristRigidPts: int = 7
