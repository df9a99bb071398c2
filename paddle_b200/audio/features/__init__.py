"""paddle.audio.features (reference: python/paddle/audio/features/layers.py): Spectrogram, MelSpectrogram,
LogMelSpectrogram, MFCC as layers over ``paddle.signal.stft``.  Input ``[N, T]`` waveforms; output
``[N, bins, frames]``."""
from __future__ import annotations

import torch

from ... import signal
from ...nn.layer.layers import Layer
from .. import functional as AF

__all__ = ["LogMelSpectrogram", "MelSpectrogram", "MFCC", "Spectrogram"]


class Spectrogram(Layer):
    """|STFT|^power.  ``win_length`` defaults to ``n_fft``, ``hop_length`` to ``win_length // 4``."""

    def __init__(self, n_fft=512, hop_length=512, win_length=None, window="hann", power=1.0, center=True,
                 pad_mode="reflect", dtype="float32"):
        super().__init__()
        assert power > 0, "Power of spectrogram must be > 0."
        self.power = power
        if win_length is None:
            win_length = n_fft
        if hop_length is None:
            hop_length = win_length // 4
        self.n_fft, self.hop_length, self.win_length, self.center, self.pad_mode = n_fft, hop_length, win_length, center, pad_mode
        self.register_buffer("fft_window", AF.get_window(window, win_length, fftbins=True, dtype=dtype), False)

    def forward(self, x):
        spec = signal.stft(x.to(self.fft_window.dtype), n_fft=self.n_fft, hop_length=self.hop_length, win_length=self.win_length,
                           window=self.fft_window, center=self.center, pad_mode=self.pad_mode)
        return spec.abs().pow(self.power)


class MelSpectrogram(Layer):
    def __init__(self, sr=22050, n_fft=2048, hop_length=512, win_length=None, window="hann", power=2.0, center=True,
                 pad_mode="reflect", n_mels=64, f_min=50.0, f_max=None, htk=False, norm="slaney", dtype="float32"):
        super().__init__()
        self._spectrogram = Spectrogram(n_fft, hop_length, win_length, window, power, center, pad_mode, dtype)
        self.n_mels, self.f_min, self.f_max, self.htk, self.norm = n_mels, f_min, f_max, htk, norm
        if f_max is None:
            f_max = sr // 2
        self.register_buffer("fbank_matrix", AF.compute_fbank_matrix(sr, n_fft, n_mels, f_min, f_max, htk, norm, dtype), False)

    def forward(self, x):
        return torch.matmul(self.fbank_matrix, self._spectrogram(x))


class LogMelSpectrogram(Layer):
    def __init__(self, sr=22050, n_fft=512, hop_length=None, win_length=None, window="hann", power=2.0, center=True,
                 pad_mode="reflect", n_mels=64, f_min=50.0, f_max=None, htk=False, norm="slaney", ref_value=1.0, amin=1e-10,
                 top_db=None, dtype="float32"):
        super().__init__()
        self._melspectrogram = MelSpectrogram(sr, n_fft, hop_length, win_length, window, power, center, pad_mode, n_mels,
                                              f_min, f_max, htk, norm, dtype)
        self.ref_value, self.amin, self.top_db = ref_value, amin, top_db

    def forward(self, x):
        return AF.power_to_db(self._melspectrogram(x), self.ref_value, self.amin, self.top_db)


class MFCC(Layer):
    def __init__(self, sr=22050, n_mfcc=40, n_fft=512, hop_length=None, win_length=None, window="hann", power=2.0,
                 center=True, pad_mode="reflect", n_mels=64, f_min=50.0, f_max=None, htk=False, norm="slaney",
                 ref_value=1.0, amin=1e-10, top_db=None, dtype="float32"):
        super().__init__()
        assert n_mfcc <= n_mels, f"n_mfcc cannot be larger than n_mels: {n_mfcc} vs {n_mels}"
        self._log_melspectrogram = LogMelSpectrogram(sr, n_fft, hop_length, win_length, window, power, center, pad_mode,
                                                     n_mels, f_min, f_max, htk, norm, ref_value, amin, top_db, dtype)
        self.register_buffer("dct_matrix", AF.create_dct(n_mfcc, n_mels, dtype=dtype), False)

    def forward(self, x):
        log_mel = self._log_melspectrogram(x)
        return torch.matmul(log_mel.transpose(1, 2), self.dct_matrix).transpose(1, 2)
