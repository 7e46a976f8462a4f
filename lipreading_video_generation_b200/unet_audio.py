"""UNetAudio -- drop-in for video-generation/diffusion/unet_audio.py:10-66.

Same constructor and forward(image, cond_image, audio, timesteps, y=None); same state_dict keys
(audio_encoder.wav2vec2.*, audio_transformer.transform.0.*, cond_conv_in.weight + the UNet torso).
The wav2vec2 encoder stays the HuggingFace module (third-party arithmetic); everything
downstream of its output -- time-mean excluded -- runs on the sm_100a kernels: Linear+ReLU projection,
the conditioning assembly (nearest resize, 1x1 cond conv, broadcast, concat) and the UNet torso.
"""
from __future__ import annotations

import torch
import torch.nn as nn

from .unet import UNetModel


class Wav2Vec2Encoder(nn.Module):
    """unet_audio.py:10-18.  `from_pretrained` needs the checkpoint in the HF cache; pass
    `pretrained=False` (or set B2D_WAV2VEC2_RANDOM_INIT=1) for a random-init wav2vec2-base."""

    def __init__(self, model_name="facebook/wav2vec2-base-960h",
                 cache_dir="/proj/vondrick/aa4870/hf-model-checkpoints", pretrained=None):
        super().__init__()
        import os
        from transformers import Wav2Vec2Config, Wav2Vec2Model
        if pretrained is None:
            pretrained = os.environ.get("B2D_WAV2VEC2_RANDOM_INIT", "0") != "1"
        if pretrained:
            self.wav2vec2 = Wav2Vec2Model.from_pretrained(model_name, cache_dir=cache_dir)
        else:
            self.wav2vec2 = Wav2Vec2Model(Wav2Vec2Config())

    def forward(self, audio_input):
        return self.wav2vec2(**audio_input, output_hidden_states=True).last_hidden_state


class AudioFeatureTransformer(nn.Module):
    """unet_audio.py:21-30 (parameter holder; the Linear+ReLU runs in the engine)."""

    def __init__(self, input_dim, output_dim):
        super().__init__()
        self.transform = nn.Sequential(nn.Linear(input_dim, output_dim), nn.ReLU())


class UNetAudio(UNetModel):
    def __init__(self, image_size, in_channels, model_channels, out_channels, num_res_blocks,
                 attention_resolutions, image_cond=True, im_cond_input_ch=3, im_cond_output_ch=64, dropout=0.1,
                 channel_mult=(1, 2, 4), conv_resample=True, dims=2, num_classes=None, use_checkpoint=False,
                 use_fp16=False, num_heads=1, num_head_channels=-1, num_heads_upsample=-1,
                 use_scale_shift_norm=False, resblock_updown=False, use_new_attention_order=False,
                 audio_feature_dim=512, projected_audio_dim=256, audio_encoder=None):
        super().__init__(image_size, in_channels + projected_audio_dim + (im_cond_output_ch if image_cond else 0),
                         model_channels, out_channels, num_res_blocks, attention_resolutions, dropout, channel_mult,
                         conv_resample, dims, num_classes, use_checkpoint, use_fp16, num_heads, num_head_channels,
                         num_heads_upsample, use_scale_shift_norm, resblock_updown, use_new_attention_order)
        self.audio_encoder = audio_encoder if audio_encoder is not None else Wav2Vec2Encoder()
        self.audio_transformer = AudioFeatureTransformer(audio_feature_dim, projected_audio_dim)
        self.projected_audio_dim = projected_audio_dim
        self.image_size = image_size
        self.image_cond = image_cond
        if self.image_cond:
            self.cond_conv_in = nn.Conv2d(in_channels=im_cond_input_ch, out_channels=im_cond_output_ch,
                                          kernel_size=1, bias=False)
        self._cached_audio = None

    def encode_audio(self, audio):
        """wav2vec2 features averaged over time: (N, audio_feature_dim).  Loop-invariant in sampling
        (test.py:57-65 recomputes it every step); `sample_images` hoists it."""
        return self.audio_encoder(audio).mean(dim=1)

    def forward(self, image, cond_image, audio, timesteps, y=None, audio_features=None):
        if image.shape[-1] != self.image_size or image.shape[-2] != self.image_size:
            # the reference's .expand(-1,-1,image_size,image_size) + cat fails for other sizes
            raise RuntimeError(f"Sizes of tensors must match: image is {tuple(image.shape[-2:])}, "
                               f"model image_size is {self.image_size}")
        feat = audio_features if audio_features is not None else self.encode_audio(audio)
        return self._torso(image, cond_image if self.image_cond else None, feat.float(), timesteps)
