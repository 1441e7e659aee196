"""Host-side mirror of the reference decoder (``models/decoder.py``).

Parameter containers with the reference's module tree / ``state_dict`` keys
(``decoder.py:41-88``).  The arithmetic (reflect-pad 3x3 implicit-GEMM convs with
folded BatchNorm + ELU, align_corners=True upsampling, skip concat, sigmoid
heads — ``decoder.py:90-173``) runs in ``engine.DepthEngine`` on the sm_100a kernels.
"""
from __future__ import annotations

import numpy as np
import torch.nn as nn

NUM_CH_DEC = (256, 128, 64, 32, 16)  # decoder.py:49


class Conv3x3(nn.Module):
    """decoder.py:8-19 — ReflectionPad2d(1) + 3x3 conv (container)."""

    def __init__(self, in_channels: int, out_channels: int):
        super().__init__()
        self.conv = nn.Sequential(nn.ReflectionPad2d(1), nn.Conv2d(int(in_channels), int(out_channels), 3))


class ConvBlock(nn.Module):
    """decoder.py:22-34 — Conv3x3 + BatchNorm2d + ELU (container)."""

    def __init__(self, in_channels: int, out_channels: int):
        super().__init__()
        self.conv = nn.Sequential(
            Conv3x3(in_channels, out_channels), nn.BatchNorm2d(int(out_channels)), nn.ELU(inplace=True)
        )


class DepthAnythingDecoder(nn.Module):
    """Mirror of ``models/decoder.py:37`` (same ctor signature, same keys)."""

    def __init__(self, num_ch_enc, scales=range(4), num_output_channels=1, use_skips=True):
        super().__init__()
        self.num_output_channels = num_output_channels
        self.use_skips = use_skips
        self.scales = scales
        self.num_ch_dec = np.array(NUM_CH_DEC)
        self.debug = False
        if not use_skips or num_output_channels != 1 or list(scales) != [0, 1, 2, 3] or list(num_ch_enc) != [256] * 4:
            raise ValueError("only the configuration the reference models instantiate is supported "
                             "(num_ch_enc=[256]*4, scales=range(4), 1 output channel, use_skips=True)")
        self.convs = nn.ModuleDict()
        self.convs["init_conv"] = ConvBlock(num_ch_enc[-1], self.num_ch_dec[0])
        for i in range(4):
            out_ch = self.num_ch_dec[i + 1]
            self.convs[f"upconv_{i}_0"] = ConvBlock(self.num_ch_dec[i], out_ch)
            in_ch = out_ch
            if i < len(num_ch_enc) - 1:
                self.convs[f"skip_{i}"] = ConvBlock(num_ch_enc[-(i + 2)], out_ch)
                in_ch *= 2
            self.convs[f"upconv_{i}_1"] = ConvBlock(in_ch, out_ch)
        self.convs["mars_adapt"] = ConvBlock(self.num_ch_dec[-1], self.num_ch_dec[-1])
        for s in self.scales:
            self.convs[f"dispconv_{s}"] = nn.Conv2d(int(self.num_ch_dec[s + 1]), num_output_channels, 3, padding=1)
