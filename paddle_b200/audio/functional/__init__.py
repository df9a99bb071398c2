"""paddle.audio.functional (reference: python/paddle/audio/functional/functional.py): mel scale, filter banks, DCT,
decibel conversion, windows."""
from __future__ import annotations

import math

import torch

from .window import get_window  # noqa: F401

__all__ = ["compute_fbank_matrix", "create_dct", "fft_frequencies", "hz_to_mel", "mel_frequencies", "mel_to_hz",
           "power_to_db", "get_window"]

_DTYPES = {"float32": torch.float32, "float64": torch.float64}

# Slaney's auditory-toolbox mel scale: linear below 1 kHz (200/3 Hz per mel), logarithmic above (27 mels per factor 6.4)
_F_SP = 200.0 / 3
_MIN_LOG_HZ = 1000.0
_MIN_LOG_MEL = _MIN_LOG_HZ / _F_SP
_LOGSTEP = math.log(6.4) / 27.0


def _dtype(dtype):
    if isinstance(dtype, torch.dtype):
        return dtype
    if dtype not in _DTYPES:
        raise ValueError(f"dtype should be float32 or float64, got {dtype}")
    return _DTYPES[dtype]


def hz_to_mel(freq, htk=False):
    """Hz -> mel.  ``htk=True``: 2595 log10(1 + f / 700); default: Slaney's piecewise scale."""
    if htk:
        if isinstance(freq, torch.Tensor):
            return 2595.0 * torch.log10(1.0 + freq / 700.0)
        return 2595.0 * math.log10(1.0 + freq / 700.0)
    if isinstance(freq, torch.Tensor):
        mels = freq / _F_SP
        log_part = _MIN_LOG_MEL + torch.log(freq.clamp_min(1e-10) / _MIN_LOG_HZ) / _LOGSTEP
        return torch.where(freq >= _MIN_LOG_HZ, log_part, mels)
    if freq >= _MIN_LOG_HZ:
        return _MIN_LOG_MEL + math.log(freq / _MIN_LOG_HZ) / _LOGSTEP
    return freq / _F_SP


def mel_to_hz(mel, htk=False):
    if htk:
        return 700.0 * (10.0 ** (mel / 2595.0) - 1.0)
    if isinstance(mel, torch.Tensor):
        freqs = _F_SP * mel
        log_part = _MIN_LOG_HZ * torch.exp(_LOGSTEP * (mel - _MIN_LOG_MEL))
        return torch.where(mel >= _MIN_LOG_MEL, log_part, freqs)
    if mel >= _MIN_LOG_MEL:
        return _MIN_LOG_HZ * math.exp(_LOGSTEP * (mel - _MIN_LOG_MEL))
    return _F_SP * mel


def mel_frequencies(n_mels=64, f_min=0.0, f_max=11025.0, htk=False, dtype="float32"):
    """``n_mels`` frequencies evenly spaced on the mel scale between ``f_min`` and ``f_max`` (in Hz)."""
    lo, hi = hz_to_mel(f_min, htk), hz_to_mel(f_max, htk)
    mels = torch.linspace(lo, hi, n_mels, dtype=_dtype(dtype))
    return mel_to_hz(mels, htk)


def fft_frequencies(sr, n_fft, dtype="float32"):
    """Centre frequencies of the ``1 + n_fft // 2`` one-sided FFT bins."""
    return torch.linspace(0, float(sr) / 2, int(1 + n_fft // 2), dtype=_dtype(dtype))


def compute_fbank_matrix(sr, n_fft, n_mels=64, f_min=0.0, f_max=None, htk=False, norm="slaney", dtype="float32"):
    """[n_mels, 1 + n_fft // 2] triangular mel filters.  ``norm="slaney"``: every triangle has unit area (divide by
    the width of its band); a number p: L-p normalise every filter; None: peak 1."""
    td = _dtype(dtype)
    if f_max is None:
        f_max = float(sr) / 2
    fftfreqs = fft_frequencies(sr, n_fft, torch.float64)
    mel_f = mel_frequencies(n_mels + 2, f_min, f_max, htk, torch.float64)
    fdiff = mel_f[1:] - mel_f[:-1]
    ramps = mel_f.unsqueeze(1) - fftfreqs.unsqueeze(0)
    lower = -ramps[:-2] / fdiff[:-1].unsqueeze(1)
    upper = ramps[2:] / fdiff[1:].unsqueeze(1)
    weights = torch.clamp_min(torch.minimum(lower, upper), 0.0)
    if norm == "slaney":
        enorm = 2.0 / (mel_f[2:n_mels + 2] - mel_f[:n_mels])
        weights = weights * enorm.unsqueeze(1)
    elif isinstance(norm, (int, float)):
        weights = torch.nn.functional.normalize(weights, p=float(norm), dim=-1)
    elif norm is not None:
        raise ValueError(f"unsupported norm {norm!r}")
    return weights.to(td)


def power_to_db(spect, ref_value=1.0, amin=1e-10, top_db=80.0):
    """10 log10(max(spect, amin)) - 10 log10(max(ref_value, amin)), floored at (max - top_db) when ``top_db`` is set."""
    if amin <= 0:
        raise Exception("amin must be strictly positive")
    if ref_value <= 0:
        raise Exception("ref_value must be strictly positive")
    log_spec = 10.0 * torch.log10(torch.clamp_min(spect, amin))
    log_spec = log_spec - 10.0 * math.log10(max(ref_value, amin))
    if top_db is not None:
        if top_db < 0:
            raise Exception("top_db must be non-negative")
        log_spec = torch.clamp_min(log_spec, float(log_spec.max()) - top_db)
    return log_spec


def create_dct(n_mfcc, n_mels, norm="ortho", dtype="float32"):
    """[n_mels, n_mfcc] DCT-II basis: mfcc = log_mel^T @ dct."""
    td = _dtype(dtype)
    n = torch.arange(n_mels, dtype=torch.float64)
    k = torch.arange(n_mfcc, dtype=torch.float64).unsqueeze(1)
    dct = torch.cos(math.pi / float(n_mels) * (n + 0.5) * k)
    if norm is None:
        dct = dct * 2.0
    else:
        if norm != "ortho":
            raise ValueError("norm should be 'ortho' or None")
        dct[0] = dct[0] * (1.0 / math.sqrt(2.0))
        dct = dct * math.sqrt(2.0 / float(n_mels))
    return dct.t().contiguous().to(td)
