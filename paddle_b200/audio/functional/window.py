"""Window functions (reference: python/paddle/audio/functional/window.py; the definitions are scipy.signal.windows').

``get_window(window, win_length, fftbins=True)``: ``window`` is a name or ``(name, parameter...)``; ``fftbins=True``
gives the periodic window used with FFTs (computed as the symmetric window of length + 1 with its last sample dropped),
``False`` the symmetric one used for filter design.
"""
from __future__ import annotations

import math

import torch

_DTYPES = {"float32": torch.float32, "float64": torch.float64, torch.float32: torch.float32, torch.float64: torch.float64}


def _extend(M, sym):
    return (M, False) if sym else (M + 1, True)


def _truncate(w, needed):
    return w[:-1] if needed else w


def _guard(M):
    if int(M) != M or M < 0:
        raise ValueError("Window length M must be a non-negative integer")
    return M <= 1


def _general_cosine(M, a, sym, dtype):
    if _guard(M):
        return torch.ones(M, dtype=dtype)
    M, trunc = _extend(M, sym)
    fac = torch.linspace(-math.pi, math.pi, M, dtype=dtype)
    w = torch.zeros(M, dtype=dtype)
    for k, ak in enumerate(a):
        w = w + ak * torch.cos(k * fac)
    return _truncate(w, trunc)


def _general_hamming(M, alpha, sym, dtype):
    return _general_cosine(M, [alpha, 1.0 - alpha], sym, dtype)


def _hamming(M, sym, dtype):
    return _general_hamming(M, 0.54, sym, dtype)


def _hann(M, sym, dtype):
    return _general_hamming(M, 0.5, sym, dtype)


def _blackman(M, sym, dtype):
    return _general_cosine(M, [0.42, 0.50, 0.08], sym, dtype)


def _cosine(M, sym, dtype):
    if _guard(M):
        return torch.ones(M, dtype=dtype)
    M, trunc = _extend(M, sym)
    return _truncate(torch.sin(math.pi / M * (torch.arange(0, M, dtype=dtype) + 0.5)), trunc)


def _triang(M, sym, dtype):
    if _guard(M):
        return torch.ones(M, dtype=dtype)
    M, trunc = _extend(M, sym)
    n = torch.arange(1, (M + 1) // 2 + 1, dtype=dtype)
    if M % 2 == 0:
        w = (2 * n - 1.0) / M
        w = torch.cat([w, w.flip(0)])
    else:
        w = 2 * n / (M + 1.0)
        w = torch.cat([w, w[:-1].flip(0)])
    return _truncate(w, trunc)


def _bohman(M, sym, dtype):
    if _guard(M):
        return torch.ones(M, dtype=dtype)
    M, trunc = _extend(M, sym)
    fac = torch.abs(torch.linspace(-1, 1, M, dtype=dtype)[1:-1])
    w = (1 - fac) * torch.cos(math.pi * fac) + 1.0 / math.pi * torch.sin(math.pi * fac)
    w = torch.cat([torch.zeros(1, dtype=dtype), w, torch.zeros(1, dtype=dtype)])
    return _truncate(w, trunc)


def _gaussian(M, std, sym, dtype):
    if _guard(M):
        return torch.ones(M, dtype=dtype)
    M, trunc = _extend(M, sym)
    n = torch.arange(0, M, dtype=dtype) - (M - 1.0) / 2.0
    return _truncate(torch.exp(-(n ** 2) / (2 * std * std)), trunc)


def _general_gaussian(M, p, sig, sym, dtype):
    if _guard(M):
        return torch.ones(M, dtype=dtype)
    M, trunc = _extend(M, sym)
    n = torch.arange(0, M, dtype=dtype) - (M - 1.0) / 2.0
    return _truncate(torch.exp(-0.5 * torch.abs(n / sig) ** (2 * p)), trunc)


def _exponential(M, center, tau, sym, dtype):
    if sym and center is not None:
        raise ValueError("If sym==True, center must be None.")
    if _guard(M):
        return torch.ones(M, dtype=dtype)
    M, trunc = _extend(M, sym)
    if center is None:
        center = (M - 1) / 2
    n = torch.arange(0, M, dtype=dtype)
    return _truncate(torch.exp(-torch.abs(n - center) / tau), trunc)


def _tukey(M, alpha, sym, dtype):
    if _guard(M):
        return torch.ones(M, dtype=dtype)
    if alpha <= 0:
        return torch.ones(M, dtype=dtype)
    if alpha >= 1.0:
        return _hann(M, sym, dtype)
    M, trunc = _extend(M, sym)
    n = torch.arange(0, M, dtype=dtype)
    width = int(math.floor(alpha * (M - 1) / 2.0))
    n1, n2, n3 = n[0:width + 1], n[width + 1:M - width - 1], n[M - width - 1:]
    w1 = 0.5 * (1 + torch.cos(math.pi * (-1 + 2.0 * n1 / alpha / (M - 1))))
    w2 = torch.ones(n2.shape, dtype=dtype)
    w3 = 0.5 * (1 + torch.cos(math.pi * (-2.0 / alpha + 1 + 2.0 * n3 / alpha / (M - 1))))
    return _truncate(torch.cat([w1, w2, w3]), trunc)


def _taylor(M, nbar, sll, norm, sym, dtype):
    """Taylor window: ``nbar`` nearly constant-level sidelobes at ``sll`` dB below the main lobe."""
    if _guard(M):
        return torch.ones(M, dtype=dtype)
    M, trunc = _extend(M, sym)
    B = 10 ** (sll / 20)
    A = math.acosh(B) / math.pi
    s2 = nbar ** 2 / (A ** 2 + (nbar - 0.5) ** 2)
    ma = torch.arange(1, nbar, dtype=dtype)
    Fm = torch.empty(nbar - 1, dtype=dtype)
    signs = torch.empty_like(ma)
    signs[::2] = 1
    signs[1::2] = -1
    m2 = ma * ma
    for mi in range(len(ma)):
        numer = signs[mi] * torch.prod(1 - m2[mi] / s2 / (A ** 2 + (ma - 0.5) ** 2))
        if mi == 0:
            denom = 2 * torch.prod(1 - m2[mi] / m2[mi + 1:])
        elif mi == len(ma) - 1:
            denom = 2 * torch.prod(1 - m2[mi] / m2[:mi])
        else:
            denom = 2 * torch.prod(1 - m2[mi] / m2[:mi]) * torch.prod(1 - m2[mi] / m2[mi + 1:])
        Fm[mi] = numer / denom

    def W(n):
        return 1 + 2 * torch.matmul(Fm.unsqueeze(0), torch.cos(2 * math.pi * ma.unsqueeze(1) * (n - M / 2.0 + 0.5) / M))
    w = W(torch.arange(0, M, dtype=dtype))
    if norm:
        w = w / W(torch.tensor((M - 1) / 2, dtype=dtype))
    return _truncate(w.squeeze(0), trunc)


# name -> (function, names of the positional parameters a tuple spec carries)
_WINDOWS = {
    "hamming": (_hamming, 0), "hann": (_hann, 0), "blackman": (_blackman, 0), "cosine": (_cosine, 0), "triang": (_triang, 0),
    "bohman": (_bohman, 0), "gaussian": (_gaussian, 1), "general_gaussian": (_general_gaussian, 2),
    "exponential": (_exponential, 2), "tukey": (_tukey, 1), "taylor": (_taylor, 3), "general_cosine": (_general_cosine, 1),
    "general_hamming": (_general_hamming, 1),
}
_DEFAULTS = {"exponential": (None, 1.0), "tukey": (0.5,), "taylor": (4, 30, True)}


def get_window(window, win_length, fftbins=True, dtype="float64"):
    sym = not fftbins
    if dtype not in _DTYPES:
        raise ValueError(f"dtype should be float32 or float64, got {dtype}")
    td = _DTYPES[dtype]
    args = ()
    if isinstance(window, tuple):
        name = window[0]
        args = tuple(window[1:])
    elif isinstance(window, str):
        name = window
        if name in ("gaussian", "general_gaussian", "general_cosine", "general_hamming"):
            raise ValueError(f"The '{name}' window needs one or more parameters -- pass a tuple.")
    else:
        raise ValueError(f"The type of window should be str or tuple, got {type(window)}")
    if name not in _WINDOWS:
        raise ValueError(f"Unknown window type: {name}")
    fn, nparams = _WINDOWS[name]
    if not args and name in _DEFAULTS:
        args = _DEFAULTS[name]
    if len(args) < nparams and name in _DEFAULTS:
        args = args + _DEFAULTS[name][len(args):]
    if len(args) != nparams:
        raise ValueError(f"window '{name}' takes {nparams} parameter(s), got {len(args)}")
    return fn(int(win_length), *args, sym, td)
