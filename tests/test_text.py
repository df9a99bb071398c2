"""paddle.text: Viterbi decoding against brute force, dataset parsers on miniature archives in the corpora's formats
(reference: python/paddle/text/{viterbi_decode.py,datasets/*.py})."""
import gzip
import io
import itertools
import os
import tarfile
import zipfile

import numpy as np
import pytest
import torch

import paddle
import paddle.text as text


def _best_by_enumeration(pot, trans, length, with_tags):
    n = pot.shape[-1]
    best = None
    for seq in itertools.product(range(n), repeat=length):
        v = sum(float(pot[i, seq[i]]) for i in range(length)) + sum(float(trans[seq[i], seq[i + 1]]) for i in range(length - 1))
        if with_tags:
            v += float(trans[-1, seq[0]]) + float(trans[seq[-1], -2])
        if best is None or v > best[0]:
            best = (v, list(seq))
    return best


@pytest.mark.parametrize("with_tags", [True, False])
def test_viterbi_decode_matches_enumeration(with_tags):
    torch.manual_seed(0)
    pot = torch.randn(5, 5, 4, dtype=torch.float64)
    trans = torch.randn(4, 4, dtype=torch.float64)
    lengths = torch.tensor([5, 3, 1, 4, 2])
    scores, paths = text.viterbi_decode(pot, trans, lengths, with_tags)
    assert scores.shape == (5,) and paths.shape == (5, 5) and paths.dtype == torch.int64
    for s in range(5):
        n = int(lengths[s])
        want_score, want_path = _best_by_enumeration(pot[s], trans, n, with_tags)
        assert abs(want_score - float(scores[s])) < 1e-9
        assert paths[s, :n].tolist() == want_path and int(paths[s, n:].sum()) == 0
    # the path is as long as the longest sequence, the layer is the same computation
    s2, p2 = text.ViterbiDecoder(trans, with_tags)(pot[:, :, :], torch.tensor([2, 3, 1, 3, 2]))
    assert p2.shape == (5, 3)
    torch.testing.assert_close(s2[2], text.viterbi_decode(pot[2:3], trans, torch.tensor([1]), with_tags)[0][0])
    with pytest.raises(ValueError):
        text.viterbi_decode(pot, trans[:3], lengths)


def _tar_add(tar, name, data):
    info = tarfile.TarInfo(name)
    info.size = len(data)
    tar.addfile(info, io.BytesIO(data))


def test_uci_housing(tmp_path):
    rng = np.random.RandomState(0)
    table = rng.rand(50, 14) * 10
    path = tmp_path / "housing.data"
    path.write_text("\n".join(" ".join(f"{v:.6f}" for v in row) for row in table))
    train, test = text.UCIHousing(str(path), "train"), text.UCIHousing(str(path), "test")
    assert len(train) == 40 and len(test) == 10
    x, y = train[3]
    assert x.shape == (13,) and y.shape == (1,) and x.dtype == np.float32
    parsed = np.loadtxt(str(path))
    want = (parsed[:, :13] - parsed[:, :13].mean(0)) / (parsed[:, :13].max(0) - parsed[:, :13].min(0))
    np.testing.assert_allclose(x, want[3], rtol=1e-5)
    np.testing.assert_allclose(test[0][1], parsed[40, 13:], rtol=1e-6)
    with pytest.raises(RuntimeError):
        text.UCIHousing()
    with pytest.raises(AssertionError):
        text.UCIHousing(str(path), mode="dev")


def test_imdb_and_imikolov(tmp_path):
    path = str(tmp_path / "aclImdb.tar.gz")
    with tarfile.open(path, "w:gz") as tar:
        _tar_add(tar, "aclImdb/train/pos/0_9.txt", b"A great, GREAT movie! great fun.\n")
        _tar_add(tar, "aclImdb/train/neg/1_2.txt", b"a bad movie; bad bad acting")
        _tar_add(tar, "aclImdb/test/pos/2_8.txt", b"great acting")
        _tar_add(tar, "aclImdb/test/neg/3_1.txt", b"bad")
        _tar_add(tar, "aclImdb/README", b"ignored")
    ds = text.Imdb(path, "train", cutoff=1)
    # counts over train + test: great 4, bad 4, a 2, movie 2, acting 2, fun 1 -> kept (> 1), by (-count, word)
    assert ds.word_idx == {b"bad": 0, b"great": 1, b"a": 2, b"acting": 3, b"movie": 4, "<unk>": 5}
    assert len(ds) == 2
    doc, label = ds[0]
    assert doc.tolist() == [2, 1, 1, 4, 1, 5] and label.tolist() == [0]            # punctuation stripped, lower-cased
    assert ds[1][1].tolist() == [1]
    assert len(text.Imdb(path, "test", cutoff=1)) == 2
    # PTB
    ptb = str(tmp_path / "simple-examples.tgz")
    with tarfile.open(ptb, "w:gz") as tar:
        _tar_add(tar, "./simple-examples/data/ptb.train.txt", b" the cat sat \n the dog sat <unk> \n")
        _tar_add(tar, "./simple-examples/data/ptb.valid.txt", b" the cat ran \n")
        _tar_add(tar, "./simple-examples/data/ptb.test.txt", b" the dog ran \n")
    ng = text.Imikolov(ptb, "NGRAM", window_size=2, mode="train", min_word_freq=0)
    # counts: <e> 3, <s> 3, the 3, cat 2, sat 2, dog 1, ran 1 ; <unk> is removed from the counts and appended last
    assert ng.word_idx == {"<e>": 0, "<s>": 1, "the": 2, "cat": 3, "sat": 4, "dog": 5, "ran": 6, "<unk>": 7}
    grams = [tuple(int(v) for v in ng[i]) for i in range(len(ng))]
    assert grams[:4] == [(1, 2), (2, 3), (3, 4), (4, 0)] and len(ng) == 4 + 5 and grams[-2:] == [(4, 7), (7, 0)]
    seq = text.Imikolov(ptb, "SEQ", window_size=-1, mode="test", min_word_freq=0)
    src, trg = seq[0]
    assert src.tolist() == [1, 2, 5, 6] and trg.tolist() == [2, 5, 6, 0]
    assert len(text.Imikolov(ptb, "SEQ", window_size=4, mode="train", min_word_freq=0)) == 1   # the 5-token source is dropped
    hi = text.Imikolov(ptb, "SEQ", mode="test", min_word_freq=2)
    assert hi.word_idx == {"<e>": 0, "<s>": 1, "the": 2, "<unk>": 3} and hi[0][0].tolist() == [1, 2, 3, 3]
    with pytest.raises(AssertionError):
        text.Imikolov(ptb, "BAG")


def test_movielens(tmp_path):
    path = str(tmp_path / "ml-1m.zip")
    with zipfile.ZipFile(path, "w") as z:
        z.writestr("ml-1m/movies.dat", "1::Toy Story (1995)::Animation|Comedy\n2::Heat (1995)::Action\n")
        z.writestr("ml-1m/users.dat", "1::F::1::10::48067\n2::M::56::16::70072\n")
        z.writestr("ml-1m/ratings.dat", "".join(f"{1 + i % 2}::{1 + (i // 2) % 2}::{1 + i % 5}::97830{i}\n" for i in range(200)))
    train = text.Movielens(path, "train", test_ratio=0.25, rand_seed=3)
    test = text.Movielens(path, "test", test_ratio=0.25, rand_seed=3)
    assert len(train) + len(test) == 200 and 20 < len(test) < 80
    item = train[0]
    assert len(item) == 8
    uid, gender, age, job, mid, cats, title, rating = item
    assert uid.tolist() in ([1], [2]) and gender.tolist() == ([1] if uid[0] == 1 else [0])
    assert age.tolist() == ([0] if uid[0] == 1 else [6]) and job.tolist() == ([10] if uid[0] == 1 else [16])
    assert cats.tolist() == ([0, 1] if mid[0] == 1 else [2]) and title.tolist() == ([0, 1] if mid[0] == 1 else [2])
    assert rating.dtype == np.float64 and float(rating[0]) in (-3.0, -1.0, 1.0, 3.0, 5.0)
    assert train.movie_info[1].title.strip() == "Toy Story" and "Toy Story" in repr(train.movie_info[1])
    assert train.user_info[2].is_male and "M" in repr(train.user_info[2])


def test_conll05_wmt14_wmt16(tmp_path):
    words = "The\ncat\nchased\nthe\nmouse\n\nDogs\nbark\n\n"
    props = ("-  (A0*   \n-  *)  \nchase  (V*)  \n-  (A1*  \n-  *)  \n\n"
             "-  (A0*)  \nbark  (V*)  \n\n")
    path = str(tmp_path / "conll05st-tests.tar.gz")
    with tarfile.open(path, "w:gz") as tar:
        _tar_add(tar, "conll05st-release/test.wsj/words/test.wsj.words.gz", gzip.compress(words.encode()))
        _tar_add(tar, "conll05st-release/test.wsj/props/test.wsj.props.gz", gzip.compress(props.encode()))
    (tmp_path / "words.dict").write_text("<unk>\nthe\ncat\nchased\nmouse\nbos\neos\nbark\n")
    (tmp_path / "verbs.dict").write_text("chase\nbark\n")
    (tmp_path / "targets.dict").write_text("B-A0\nI-A0\nB-A1\nI-A1\nB-V\nI-V\nO\n")
    ds = text.Conll05st(path, str(tmp_path / "words.dict"), str(tmp_path / "verbs.dict"), str(tmp_path / "targets.dict"), emb_file="emb")
    assert len(ds) == 2
    assert ds.label_dict == {"B-A0": 0, "I-A0": 1, "B-A1": 2, "I-A1": 3, "B-V": 4, "I-V": 5, "O": 6}
    word, n2, n1, c0, p1, p2, pred, mark, label = ds[0]
    assert word.tolist() == [0, 2, 3, 1, 4]                       # "The" is unknown (case sensitive) -> 0
    assert label.tolist() == [0, 1, 4, 2, 3] and mark.tolist() == [1, 1, 1, 1, 1] and pred.tolist() == [0] * 5
    assert n2.tolist() == [0] * 5 and n1.tolist() == [2] * 5 and c0.tolist() == [3] * 5 and p1.tolist() == [1] * 5 and p2.tolist() == [4] * 5
    word, n2, n1, c0, p1, p2, pred, mark, label = ds[1]
    assert label.tolist() == [0, 4] and mark.tolist() == [1, 1] and n2.tolist() == [5, 5] and p1.tolist() == [6, 6] and pred.tolist() == [1, 1]
    assert ds.get_dict()[1] == {"chase": 0, "bark": 1} and ds.get_embedding() == "emb"
    # WMT14
    p14 = str(tmp_path / "wmt14.tgz")
    with tarfile.open(p14, "w:gz") as tar:
        _tar_add(tar, "wmt14/src.dict", b"<s>\n<e>\n<unk>\nhello\nworld\nextra\n")
        _tar_add(tar, "wmt14/trg.dict", b"<s>\n<e>\n<unk>\nbonjour\nmonde\n")
        _tar_add(tar, "wmt14/train/train", b"hello world\tbonjour monde\nhello there\tbonjour\nbroken line\n")
        _tar_add(tar, "wmt14/test/test", b"world\tmonde\n")
    w14 = text.WMT14(p14, "train", dict_size=5)
    assert len(w14) == 2 and len(w14.get_dict()[0]) == 5
    src, trg, nxt = w14[0]
    assert src.tolist() == [0, 3, 4, 1] and trg.tolist() == [0, 3, 4] and nxt.tolist() == [3, 4, 1]
    assert w14[1][0].tolist() == [0, 3, 2, 1]
    assert w14.get_dict(reverse=True)[1][3] == "bonjour" and len(text.WMT14(p14, "test", dict_size=5)) == 1
    with pytest.raises(AssertionError):
        text.WMT14(p14, "train")
    # WMT16
    p16 = str(tmp_path / "wmt16.tar.gz")
    with tarfile.open(p16, "w:gz") as tar:
        _tar_add(tar, "wmt16/train", b"a dog runs\tein hund rennt\na dog\tein hund\n")
        _tar_add(tar, "wmt16/val", b"a cat\teine katze\n")
        _tar_add(tar, "wmt16/test", b"dog\thund\n")
    w16 = text.WMT16(p16, "train", src_dict_size=5, trg_dict_size=10, lang="en")
    assert w16.get_dict("en") == {"<s>": 0, "<e>": 1, "<unk>": 2, "a": 3, "dog": 4}
    assert w16.get_dict("de")["hund"] in (3, 4) and len(w16.get_dict("de")) == 6
    src, trg, nxt = w16[0]
    assert src.tolist() == [0, 3, 4, 2, 1] and trg[0] == 0 and nxt[-1] == 1 and trg[1:].tolist() == nxt[:-1].tolist()
    rev = text.WMT16(p16, "val", src_dict_size=10, trg_dict_size=10, lang="de")
    assert rev[0][0].tolist() == [0, 2, 2, 1] and rev[0][1].tolist() == [0, rev.get_dict("en")["a"], 2]
    assert w16.get_dict("en", reverse=True)[4] == "dog"
