"""paddle.text.datasets (reference: python/paddle/text/datasets/*.py): parsers for the classic text corpora.

The reference downloads the archives; there is no network here, so every class reads the archive given as
``data_file`` (same file the reference would have downloaded, same parsing, same item layout) and says so when it is
missing.  Vocabularies are built in memory.
"""
from __future__ import annotations

import collections
import gzip
import os
import re
import string
import tarfile
import zipfile

import numpy as np

from ...io import Dataset

__all__ = ["Conll05st", "Imdb", "Imikolov", "Movielens", "UCIHousing", "WMT14", "WMT16"]


def _need(data_file, what, download=True):
    if data_file is None:
        raise RuntimeError(f"{what}: data_file is not set and this environment has no network access to download it; pass "
                           f"data_file=<path to the archive>")
    if not os.path.exists(data_file):
        raise FileNotFoundError(f"{what}: {data_file} does not exist")
    return data_file


def _mode(mode, allowed):
    mode = mode.lower()
    assert mode in allowed, f"mode should be one of {allowed}, but got {mode}"
    return mode


# ---------------------------------------------------------------------------------------------------------- regression
class UCIHousing(Dataset):
    """Boston housing: 13 features scaled by (x - mean) / (max - min) over the whole file, price as target; the first
    80 % of the rows are the training split.  Items: (float[13], float[1])."""

    def __init__(self, data_file=None, mode="train", download=True):
        self.mode = _mode(mode, ("train", "test"))
        self.data_file = _need(data_file, "UCIHousing")
        import torch
        self.dtype = {torch.float32: "float32", torch.float64: "float64"}.get(torch.get_default_dtype(), "float32")
        self._load_data()

    def _load_data(self, feature_num=14, ratio=0.8):
        data = np.fromfile(self.data_file, sep=" ")
        data = data.reshape(data.shape[0] // feature_num, feature_num)
        span = data.max(axis=0) - data.min(axis=0)
        mean = data.mean(axis=0)
        data[:, :-1] = (data[:, :-1] - mean[:-1]) / span[:-1]
        cut = int(data.shape[0] * ratio)
        self.data = data[:cut] if self.mode == "train" else data[cut:]

    def __getitem__(self, idx):
        row = self.data[idx]
        return row[:-1].astype(self.dtype), row[-1:].astype(self.dtype)

    def __len__(self):
        return len(self.data)


# ----------------------------------------------------------------------------------------------------------- sentiment
class Imdb(Dataset):
    """aclImdb tarball.  Vocabulary: words seen more than ``cutoff`` times over train + test, most frequent first,
    ``<unk>`` last.  Items: (int64[len], int64[1]) with label 0 = positive, 1 = negative."""

    def __init__(self, data_file=None, mode="train", cutoff=150, download=True):
        self.mode = _mode(mode, ("train", "test"))
        self.data_file = _need(data_file, "Imdb")
        self.word_idx = self._build_word_dict(cutoff)
        self._load_anno()

    def _tokenize(self, pattern):
        strip = string.punctuation.encode("latin-1")
        docs = []
        with tarfile.open(self.data_file) as tarf:
            for member in tarf:
                if pattern.match(member.name):
                    docs.append(tarf.extractfile(member).read().rstrip(b"\n\r").translate(None, strip).lower().split())
        return docs

    def _build_word_dict(self, cutoff):
        freq = collections.Counter()
        for doc in self._tokenize(re.compile(r"aclImdb/((train)|(test))/((pos)|(neg))/.*\.txt$")):
            freq.update(doc)
        kept = sorted(((w, c) for w, c in freq.items() if c > cutoff), key=lambda x: (-x[1], x[0]))
        word_idx = {w: i for i, (w, _) in enumerate(kept)}
        word_idx["<unk>"] = len(kept)
        return word_idx

    def _load_anno(self):
        unk = self.word_idx["<unk>"]
        self.docs, self.labels = [], []
        for label, kind in enumerate(("pos", "neg")):
            for doc in self._tokenize(re.compile(rf"aclImdb/{self.mode}/{kind}/.*\.txt$")):
                self.docs.append([self.word_idx.get(w, unk) for w in doc])
                self.labels.append(label)

    def __getitem__(self, idx):
        return np.array(self.docs[idx]), np.array([self.labels[idx]])

    def __len__(self):
        return len(self.docs)


# ------------------------------------------------------------------------------------------------------ language model
class Imikolov(Dataset):
    """Penn Treebank (Mikolov's simple-examples tarball).  ``data_type="NGRAM"``: every window of ``window_size`` ids
    over ``<s> sentence <e>``; ``"SEQ"``: (``<s>`` + ids, ids + ``<e>``), sentences longer than ``window_size`` dropped
    when it is positive.  Vocabulary: words with count > ``min_word_freq`` over train + valid, ``<unk>`` last."""

    TRAIN = "./simple-examples/data/ptb.train.txt"
    VALID = "./simple-examples/data/ptb.valid.txt"

    def __init__(self, data_file=None, data_type="NGRAM", window_size=-1, mode="train", min_word_freq=50, download=True):
        assert data_type.upper() in ("NGRAM", "SEQ"), f"data type should be 'NGRAM', 'SEQ', but got {data_type}"
        self.data_type = data_type.upper()
        self.mode = _mode(mode, ("train", "test"))
        self.window_size, self.min_word_freq = window_size, min_word_freq
        self.data_file = _need(data_file, "Imikolov")
        self.word_idx = self._build_word_dict()
        self._load_anno()

    @staticmethod
    def _member(tf, name):
        for candidate in (name, name[2:]):
            try:
                return tf.extractfile(candidate)
            except KeyError:
                continue
        raise KeyError(name)

    def _build_word_dict(self):
        freq = collections.Counter()
        with tarfile.open(self.data_file) as tf:
            for name in (self.TRAIN, self.VALID):
                for line in self._member(tf, name):
                    freq.update(line.decode().strip().split())
                    freq["<s>"] += 1
                    freq["<e>"] += 1
        freq.pop("<unk>", None)
        kept = sorted(((w, c) for w, c in freq.items() if c > self.min_word_freq), key=lambda x: (-x[1], x[0]))
        word_idx = {w: i for i, (w, _) in enumerate(kept)}
        word_idx["<unk>"] = len(kept)
        return word_idx

    def _load_anno(self):
        self.data = []
        unk = self.word_idx["<unk>"]
        name = self.TRAIN if self.mode == "train" else "./simple-examples/data/ptb.test.txt"
        with tarfile.open(self.data_file) as tf:
            for line in self._member(tf, name):
                words = line.decode().strip().split()
                if self.data_type == "NGRAM":
                    assert self.window_size > -1, "Invalid gram length"
                    ids = [self.word_idx.get(w, unk) for w in ["<s>"] + words + ["<e>"]]
                    for i in range(self.window_size, len(ids) + 1):
                        self.data.append(tuple(ids[i - self.window_size:i]))
                else:
                    ids = [self.word_idx.get(w, unk) for w in words]
                    src, trg = [self.word_idx["<s>"]] + ids, ids + [self.word_idx["<e>"]]
                    if self.window_size > 0 and len(src) > self.window_size:
                        continue
                    self.data.append((src, trg))

    def __getitem__(self, idx):
        return tuple(np.array(d) for d in self.data[idx])

    def __len__(self):
        return len(self.data)


# --------------------------------------------------------------------------------------------------------- recommender
age_table = [1, 18, 25, 35, 45, 50, 56]


class MovieInfo:
    def __init__(self, index, categories, title):
        self.index, self.categories, self.title = int(index), categories, title

    def value(self, categories_dict, movie_title_dict):
        return [[self.index], [categories_dict[c] for c in self.categories],
                [movie_title_dict[w.lower()] for w in self.title.split()]]

    def __repr__(self):
        return f"<MovieInfo id({self.index}), title({self.title}), categories({self.categories})>"


class UserInfo:
    def __init__(self, index, gender, age, job_id):
        self.index, self.is_male = int(index), gender == "M"
        self.age, self.job_id = age_table.index(int(age)), int(job_id)

    def value(self):
        return [[self.index], [0 if self.is_male else 1], [self.age], [self.job_id]]

    def __repr__(self):
        return f"<UserInfo id({self.index}), gender({'M' if self.is_male else 'F'}), age({age_table[self.age]}), job({self.job_id})>"


class Movielens(Dataset):
    """MovieLens 1M zip.  Items: (user id, gender, age bucket, job, movie id, category ids, title word ids, rating) with
    the rating rescaled to [-3, 5] by ``2 r - 5``; a rating goes to the test split with probability ``test_ratio``
    (numpy generator seeded with ``rand_seed``)."""

    def __init__(self, data_file=None, mode="train", test_ratio=0.1, rand_seed=0, download=True):
        self.mode = _mode(mode, ("train", "test"))
        self.data_file = _need(data_file, "Movielens")
        self.test_ratio, self.rand_seed = test_ratio, rand_seed
        np.random.seed(rand_seed)
        self._load_meta_info()
        self._load_data()

    def _load_meta_info(self):
        title_year = re.compile(r"^(.*)\((\d+)\)$")
        self.movie_info, self.user_info = {}, {}
        words, categories = [], []
        with zipfile.ZipFile(self.data_file) as package:
            with package.open("ml-1m/movies.dat") as f:
                for line in f:
                    movie_id, title, cats = line.decode("latin").strip().split("::")
                    cats = cats.split("|")
                    title = title_year.match(title).group(1)
                    self.movie_info[int(movie_id)] = MovieInfo(movie_id, cats, title)
                    categories += [c for c in cats if c not in categories]
                    words += [w.lower() for w in title.split()]
            with package.open("ml-1m/users.dat") as f:
                for line in f:
                    uid, gender, age, job, _ = line.decode("latin").strip().split("::")
                    self.user_info[int(uid)] = UserInfo(uid, gender, age, job)
        self.movie_title_dict = {w: i for i, w in enumerate(dict.fromkeys(words))}
        self.categories_dict = {c: i for i, c in enumerate(categories)}

    def _load_data(self):
        self.data = []
        is_test = self.mode == "test"
        with zipfile.ZipFile(self.data_file) as package:
            with package.open("ml-1m/ratings.dat") as f:
                for line in f:
                    if (np.random.random() < self.test_ratio) == is_test:
                        uid, mov_id, rating, _ = line.decode("latin").strip().split("::")
                        usr, mov = self.user_info[int(uid)], self.movie_info[int(mov_id)]
                        self.data.append(usr.value() + mov.value(self.categories_dict, self.movie_title_dict)
                                         + [[float(rating) * 2 - 5.0]])

    def __getitem__(self, idx):
        return tuple(np.array(d) for d in self.data[idx])

    def __len__(self):
        return len(self.data)


# ------------------------------------------------------------------------------------------------ semantic role labels
def _props_to_bio(column):
    """One predicate's bracket column of a CoNLL-05 props file -> BIO tags: ``(A0*`` opens A0, ``*)`` closes it, ``*``
    continues what is open (or is outside), ``(V*)`` is a one-token span."""
    tags, open_tag = [], None
    for cell in column:
        if cell == "*":
            tags.append("I-" + open_tag if open_tag else "O")
        elif cell == "*)":
            tags.append("I-" + open_tag)
            open_tag = None
        elif "(" in cell:
            tag = cell[1:cell.find("*")]
            tags.append("B-" + tag)
            open_tag = None if ")" in cell else tag
        else:
            raise RuntimeError(f"Unexpected label: {cell}")
    return tags


class Conll05st(Dataset):
    """CoNLL-2005 shared task test split (the only part that is public): one item per (sentence, predicate).  Item =
    (word ids, five context-word id columns (verb -2 .. +2, repeated over the sentence), predicate id column,
    context mark, label ids).  Dictionaries come from the three dict files; unknown words map to 0.

    The label dictionary is numbered over the SORTED tag set (B-x, I-x pairs, then O) - the reference iterates a Python
    set there, so its numbering changes from run to run."""

    UNK_IDX = 0
    WORDS = "conll05st-release/test.wsj/words/test.wsj.words.gz"
    PROPS = "conll05st-release/test.wsj/props/test.wsj.props.gz"

    def __init__(self, data_file=None, word_dict_file=None, verb_dict_file=None, target_dict_file=None, emb_file=None,
                 download=True):
        self.data_file = _need(data_file, "Conll05st")
        self.word_dict_file = _need(word_dict_file, "Conll05st word_dict_file")
        self.verb_dict_file = _need(verb_dict_file, "Conll05st verb_dict_file")
        self.target_dict_file = _need(target_dict_file, "Conll05st target_dict_file")
        self.emb_file = emb_file
        self.word_dict = self._load_dict(self.word_dict_file)
        self.predicate_dict = self._load_dict(self.verb_dict_file)
        self.label_dict = self._load_label_dict(self.target_dict_file)
        self._load_anno()

    @staticmethod
    def _load_dict(filename):
        with open(filename) as f:
            return {line.strip(): i for i, line in enumerate(f)}

    @staticmethod
    def _load_label_dict(filename):
        with open(filename) as f:
            tags = sorted({line.strip()[2:] for line in f if line.strip()[:2] in ("B-", "I-")})
        d = {}
        for tag in tags:
            d["B-" + tag] = len(d)
            d["I-" + tag] = len(d)
        d["O"] = len(d)
        return d

    def _load_anno(self):
        self.sentences, self.predicates, self.labels = [], [], []
        with tarfile.open(self.data_file) as tf:
            with gzip.GzipFile(fileobj=tf.extractfile(self.WORDS)) as wf, gzip.GzipFile(fileobj=tf.extractfile(self.PROPS)) as pf:
                words, rows = [], []
                for word, props in zip(wf, pf):
                    cells = props.decode().strip().split()
                    if cells:
                        words.append(word.decode().strip())
                        rows.append(cells)
                        continue
                    if rows:                                         # blank line: the sentence is complete
                        columns = list(zip(*rows))
                        verbs = [v for v in columns[0] if v != "-"]
                        for verb, column in zip(verbs, columns[1:]):
                            self.sentences.append(words)
                            self.predicates.append(verb)
                            self.labels.append(_props_to_bio(column))
                    words, rows = [], []

    def __getitem__(self, idx):
        sentence, predicate, labels = self.sentences[idx], self.predicates[idx], self.labels[idx]
        n = len(sentence)
        v = labels.index("B-V")
        mark = [0] * n
        context = []
        for off in (-2, -1, 0, 1, 2):
            j = v + off
            if 0 <= j < n:
                mark[j] = 1
                context.append(sentence[j])
            else:
                context.append("bos" if off < 0 else "eos")
        get = self.word_dict.get
        columns = [[get(w, self.UNK_IDX) for w in sentence]] + [[get(c, self.UNK_IDX)] * n for c in context]
        columns.append([self.predicate_dict.get(predicate)] * n)
        columns.append(mark)
        columns.append([self.label_dict.get(t) for t in labels])
        return tuple(np.array(c) for c in columns)

    def __len__(self):
        return len(self.sentences)

    def get_dict(self):
        return self.word_dict, self.predicate_dict, self.label_dict

    def get_embedding(self):
        return self.emb_file


# --------------------------------------------------------------------------------------------------------- translation
START, END, UNK = "<s>", "<e>", "<unk>"


class WMT14(Dataset):
    """The pre-tokenised WMT14 en-fr subset tarball: ``*/src.dict``, ``*/trg.dict`` (first ``dict_size`` lines are the
    vocabularies; lines 0..2 are <s>, <e>, <unk>) and ``<mode>/<mode>`` files of ``source \\t target`` lines.  Items:
    (<s> src <e>, <s> trg, trg <e>); pairs with a side longer than 80 are dropped."""

    UNK_IDX = 2

    def __init__(self, data_file=None, mode="train", dict_size=-1, download=True):
        self.mode = _mode(mode, ("train", "test", "gen"))
        self.data_file = _need(data_file, "WMT14")
        assert dict_size > 0, "dict_size should be set as positive number"
        self.dict_size = dict_size
        self._load_data()

    def _load_data(self):
        def read_dict(fd):
            return {line.strip().decode(): i for i, line in zip(range(self.dict_size), fd)}
        self.src_ids, self.trg_ids, self.trg_ids_next = [], [], []
        with tarfile.open(self.data_file, mode="r") as f:
            names = f.getnames()
            src = [n for n in names if n.endswith("src.dict")]
            trg = [n for n in names if n.endswith("trg.dict")]
            assert len(src) == 1 and len(trg) == 1, "the archive should hold exactly one src.dict and one trg.dict"
            self.src_dict, self.trg_dict = read_dict(f.extractfile(src[0])), read_dict(f.extractfile(trg[0]))
            for name in (n for n in names if n.endswith(f"{self.mode}/{self.mode}")):
                for line in f.extractfile(name):
                    parts = line.decode().strip().split("\t")
                    if len(parts) != 2:
                        continue
                    src_ids = [self.src_dict.get(w, self.UNK_IDX) for w in [START] + parts[0].split() + [END]]
                    trg_ids = [self.trg_dict.get(w, self.UNK_IDX) for w in parts[1].split()]
                    if len(src_ids) > 80 or len(trg_ids) > 80:
                        continue
                    self.src_ids.append(src_ids)
                    self.trg_ids.append([self.trg_dict[START]] + trg_ids)
                    self.trg_ids_next.append(trg_ids + [self.trg_dict[END]])

    def __getitem__(self, idx):
        return np.array(self.src_ids[idx]), np.array(self.trg_ids[idx]), np.array(self.trg_ids_next[idx])

    def __len__(self):
        return len(self.src_ids)

    def get_dict(self, reverse=False):
        if reverse:
            return {v: k for k, v in self.src_dict.items()}, {v: k for k, v in self.trg_dict.items()}
        return self.src_dict, self.trg_dict


class WMT16(Dataset):
    """Multi30k en-de tarball (``wmt16/train``, ``wmt16/val``, ``wmt16/test``: ``english \\t german`` lines).  The
    vocabularies are built from the training file: <s>, <e>, <unk>, then words by falling frequency, ``*_dict_size``
    entries in all.  ``lang`` is the SOURCE language.  Items: (<s> src <e>, <s> trg, trg <e>)."""

    def __init__(self, data_file=None, mode="train", src_dict_size=-1, trg_dict_size=-1, lang="en", download=True):
        self.mode = _mode(mode, ("train", "test", "val"))
        self.data_file = _need(data_file, "WMT16")
        assert lang in ("en", "de"), "lang should be 'en' or 'de'"
        assert src_dict_size > 0 and trg_dict_size > 0, "dict_size should be set as positive number"
        self.lang = lang
        self.src_dict_size, self.trg_dict_size = src_dict_size, trg_dict_size
        self._dicts = {lang: self._build_dict(lang, src_dict_size),
                       ("de" if lang == "en" else "en"): self._build_dict("de" if lang == "en" else "en", trg_dict_size)}
        self.src_dict = self._dicts[lang]
        self.trg_dict = self._dicts["de" if lang == "en" else "en"]
        self.src_dict_size, self.trg_dict_size = len(self.src_dict), len(self.trg_dict)
        self._load_data()

    def _build_dict(self, lang, dict_size):
        col = 0 if lang == "en" else 1
        freq = collections.Counter()
        with tarfile.open(self.data_file, mode="r") as f:
            for line in f.extractfile("wmt16/train"):
                parts = line.decode().strip().split("\t")
                if len(parts) == 2:
                    freq.update(parts[col].split())
        words = [START, END, UNK] + [w for w, _ in sorted(freq.items(), key=lambda x: -x[1])][:max(dict_size - 3, 0)]
        return {w: i for i, w in enumerate(words)}

    def _load_data(self):
        start, end, unk = self.src_dict[START], self.src_dict[END], self.src_dict[UNK]
        src_col = 0 if self.lang == "en" else 1
        self.src_ids, self.trg_ids, self.trg_ids_next = [], [], []
        with tarfile.open(self.data_file, mode="r") as f:
            for line in f.extractfile(f"wmt16/{self.mode}"):
                parts = line.decode().strip().split("\t")
                if len(parts) != 2:
                    continue
                trg_ids = [self.trg_dict.get(w, unk) for w in parts[1 - src_col].split()]
                self.src_ids.append([start] + [self.src_dict.get(w, unk) for w in parts[src_col].split()] + [end])
                self.trg_ids.append([start] + trg_ids)
                self.trg_ids_next.append(trg_ids + [end])

    def __getitem__(self, idx):
        return np.array(self.src_ids[idx]), np.array(self.trg_ids[idx]), np.array(self.trg_ids_next[idx])

    def __len__(self):
        return len(self.src_ids)

    def get_dict(self, lang, reverse=False):
        d = self._dicts[lang]
        return {v: k for k, v in d.items()} if reverse else d
