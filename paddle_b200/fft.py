"""paddle.fft (reference: python/paddle/fft.py) on torch.fft; ``axis``/``axes`` argument names kept."""
from __future__ import annotations

import torch


def _1d(fn):
    def f(x, n=None, axis=-1, norm="backward", name=None):
        return fn(x, n=n, dim=axis, norm=norm)
    f.__name__ = fn.__name__
    return f


def _nd(fn):
    def f(x, s=None, axes=None, norm="backward", name=None):
        return fn(x, s=s, dim=axes, norm=norm)
    f.__name__ = fn.__name__
    return f


def _2d(fn):
    def f(x, s=None, axes=(-2, -1), norm="backward", name=None):
        return fn(x, s=s, dim=axes, norm=norm)
    f.__name__ = fn.__name__
    return f


fft, ifft, rfft, irfft, hfft, ihfft = (_1d(getattr(torch.fft, n)) for n in ("fft", "ifft", "rfft", "irfft", "hfft", "ihfft"))
fftn, ifftn, rfftn, irfftn = (_nd(getattr(torch.fft, n)) for n in ("fftn", "ifftn", "rfftn", "irfftn"))
fft2, ifft2, rfft2, irfft2 = (_2d(getattr(torch.fft, n)) for n in ("fft2", "ifft2", "rfft2", "irfft2"))


def fftfreq(n, d=1.0, dtype=None, name=None):
    return torch.fft.fftfreq(n, d=d, dtype=dtype)


def rfftfreq(n, d=1.0, dtype=None, name=None):
    return torch.fft.rfftfreq(n, d=d, dtype=dtype)


def fftshift(x, axes=None, name=None):
    return torch.fft.fftshift(x, dim=axes)


def ifftshift(x, axes=None, name=None):
    return torch.fft.ifftshift(x, dim=axes)


# Hermitian-input transforms in 2 / n dimensions.  hfftn(x) = irfftn(conj(x)) with the normalisation roles swapped
# ("backward" <-> "forward"), ihfftn(x) = conj(rfftn(x)) likewise: a Hermitian-symmetric signal has a real spectrum.
_SWAP = {"backward": "forward", "forward": "backward", "ortho": "ortho"}


def hfftn(x, s=None, axes=None, norm="backward", name=None):
    return torch.fft.irfftn(torch.conj(x), s=s, dim=axes, norm=_SWAP[norm])


def ihfftn(x, s=None, axes=None, norm="backward", name=None):
    return torch.conj(torch.fft.rfftn(x, s=s, dim=axes, norm=_SWAP[norm]))


def hfft2(x, s=None, axes=(-2, -1), norm="backward", name=None):
    return hfftn(x, s, axes, norm)


def ihfft2(x, s=None, axes=(-2, -1), norm="backward", name=None):
    return ihfftn(x, s, axes, norm)
