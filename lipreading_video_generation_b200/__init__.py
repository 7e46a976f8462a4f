"""B200-native denoiser train step / sampler step of wdas03/lipreading-video-generation's
video-generation/diffusion, behind the reference's Python API."""
from .loss import MSELoss, mse_loss
from .schedulers import CosineNoiseScheduler, LinearNoiseScheduler, LinearNoiseSchedulerV2
from .unet import UNetModel
from .unet_audio import AudioFeatureTransformer, UNetAudio, Wav2Vec2Encoder
from .sampling import GraphedDenoiser, sample_images
from .optim import Adam
from .dataset import DeviceVideoStore, FrameItem, TalkingFaceFrameBatcher

__all__ = ["UNetModel", "UNetAudio", "Wav2Vec2Encoder", "AudioFeatureTransformer", "LinearNoiseScheduler",
           "LinearNoiseSchedulerV2", "CosineNoiseScheduler", "MSELoss", "mse_loss", "sample_images", "GraphedDenoiser", "Adam", "DeviceVideoStore", "FrameItem",
           "TalkingFaceFrameBatcher"]
