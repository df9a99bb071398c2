"""paddle.text (reference: python/paddle/text/__init__.py)."""
from . import datasets  # noqa: F401
from .datasets import Conll05st, Imdb, Imikolov, Movielens, UCIHousing, WMT14, WMT16  # noqa: F401
from .viterbi_decode import ViterbiDecoder, viterbi_decode  # noqa: F401

__all__ = ["Conll05st", "Imdb", "Imikolov", "Movielens", "UCIHousing", "WMT14", "WMT16", "ViterbiDecoder", "viterbi_decode"]
