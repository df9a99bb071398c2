"""paddle.audio against scipy restatements (reference: python/paddle/audio/{functional,features,backends,datasets})."""
import math
import os
import wave

import numpy as np
import pytest
import scipy.fft
import scipy.signal
import scipy.signal.windows as W
import torch

import paddle  # noqa: F401
import paddle.audio as A
import paddle.audio.functional as AF


def test_windows_match_scipy():
    plain = {"hann": W.hann, "hamming": W.hamming, "blackman": W.blackman, "cosine": W.cosine, "triang": W.triang, "bohman": W.bohman}
    for name, ref in plain.items():
        for n in (1, 8, 9):
            for fftbins in (True, False):
                np.testing.assert_allclose(AF.get_window(name, n, fftbins).numpy(), ref(n, sym=not fftbins), atol=1e-12, err_msg=name)
    params = [(("gaussian", 2.5), lambda n, s: W.gaussian(n, 2.5, s)),
              (("general_gaussian", 1.5, 3.0), lambda n, s: W.general_gaussian(n, 1.5, 3.0, s)),
              (("exponential", None, 2.0), lambda n, s: W.exponential(n, None, 2.0, s)),
              (("tukey", 0.3), lambda n, s: W.tukey(n, 0.3, s)),
              (("taylor", 4, 30, True), lambda n, s: W.taylor(n, 4, 30, True, s)),
              (("general_hamming", 0.6), lambda n, s: W.general_hamming(n, 0.6, s)),
              (("general_cosine", [0.4, 0.5, 0.1]), lambda n, s: W.general_cosine(n, [0.4, 0.5, 0.1], s))]
    for spec, ref in params:
        for n in (8, 9):
            for fftbins in (True, False):
                np.testing.assert_allclose(AF.get_window(spec, n, fftbins).numpy(), ref(n, not fftbins), atol=1e-10, err_msg=str(spec))
    assert AF.get_window("hann", 16).dtype == torch.float64 and AF.get_window("hann", 16, dtype="float32").dtype == torch.float32
    np.testing.assert_allclose(AF.get_window("tukey", 12).numpy(), W.tukey(12, 0.5, False), atol=1e-12)      # default alpha
    for bad in ("kaiser_bessel", ("hann", 1.0), "gaussian", 3):
        with pytest.raises(ValueError):
            AF.get_window(bad, 8)


def test_mel_scale_filterbank_db_dct():
    # Slaney scale: linear to 1 kHz (200/3 Hz per mel), log above; HTK: 2595 log10(1 + f / 700)
    assert AF.hz_to_mel(500.0) == pytest.approx(7.5) and AF.hz_to_mel(1000.0) == pytest.approx(15.0)
    assert AF.hz_to_mel(6400.0) == pytest.approx(15.0 + 27.0)
    assert AF.hz_to_mel(700.0, htk=True) == pytest.approx(2595 * math.log10(2.0))
    f = torch.tensor([50.0, 440.0, 1000.0, 4000.0, 11025.0])
    for htk in (False, True):
        torch.testing.assert_close(AF.mel_to_hz(AF.hz_to_mel(f, htk), htk), f, rtol=1e-5, atol=1e-3)
        assert AF.mel_to_hz(AF.hz_to_mel(3000.0, htk), htk) == pytest.approx(3000.0)
    mf = AF.mel_frequencies(10, 0.0, 8000.0)
    assert mf.shape == (10,) and float(mf[0]) == 0.0 and float(mf[-1]) == pytest.approx(8000.0, rel=1e-5)
    mels = AF.hz_to_mel(mf.double())
    torch.testing.assert_close(mels[1:] - mels[:-1], torch.full((9,), float(mels[-1] / 9), dtype=torch.float64), rtol=1e-4, atol=1e-4)
    torch.testing.assert_close(AF.fft_frequencies(16000, 8), torch.tensor([0.0, 2000.0, 4000.0, 6000.0, 8000.0]))
    # filter bank: triangles between neighbouring mel points, unit area under "slaney"
    sr, n_fft, n_mels = 16000, 512, 20
    fb = AF.compute_fbank_matrix(sr, n_fft, n_mels, dtype="float64")
    assert fb.shape == (n_mels, 257) and float(fb.min()) >= 0
    pts = AF.mel_frequencies(n_mels + 2, 0.0, sr / 2, dtype="float64").numpy()
    freqs = np.linspace(0, sr / 2, 257)
    for i in (0, 7, 19):
        lo, c, hi = pts[i], pts[i + 1], pts[i + 2]
        tri = np.maximum(0, np.minimum((freqs - lo) / (c - lo), (hi - freqs) / (hi - c))) * 2.0 / (hi - lo)
        np.testing.assert_allclose(fb[i].numpy(), tri, atol=1e-10)
    peak1 = AF.compute_fbank_matrix(sr, n_fft, n_mels, norm=None, dtype="float64")
    assert float(peak1.max()) <= 1.0 and float(peak1.max()) > 0.8
    l1 = AF.compute_fbank_matrix(sr, n_fft, n_mels, norm=1, dtype="float64")
    torch.testing.assert_close(l1.sum(-1), torch.ones(n_mels, dtype=torch.float64))
    assert AF.compute_fbank_matrix(sr, n_fft, n_mels, f_min=100.0, f_max=4000.0, htk=True).dtype == torch.float32
    # decibels
    s = torch.tensor([1e-12, 1.0, 100.0])
    torch.testing.assert_close(AF.power_to_db(s, top_db=None), torch.tensor([-100.0, 0.0, 20.0]))
    torch.testing.assert_close(AF.power_to_db(s), torch.tensor([-60.0, 0.0, 20.0]))                # floored at max - 80
    torch.testing.assert_close(AF.power_to_db(s, ref_value=10.0, top_db=None), torch.tensor([-110.0, -10.0, 10.0]))
    with pytest.raises(Exception):
        AF.power_to_db(s, amin=0.0)
    # DCT-II, orthonormal: same as scipy
    x = np.random.RandomState(0).randn(40)
    dct = AF.create_dct(13, 40, dtype="float64")
    assert dct.shape == (40, 13)
    np.testing.assert_allclose(x @ dct.numpy(), scipy.fft.dct(x, type=2, norm="ortho")[:13], atol=1e-10)
    np.testing.assert_allclose(x @ AF.create_dct(13, 40, norm=None, dtype="float64").numpy(), scipy.fft.dct(x, type=2)[:13], atol=1e-10)


def test_feature_layers():
    torch.manual_seed(0)
    sr = 16000
    x = torch.randn(2, 4000, dtype=torch.float64)
    spec = A.features.Spectrogram(n_fft=256, hop_length=64, power=2.0, dtype="float64")(x)
    assert spec.shape == (2, 129, 4000 // 64 + 1)
    # against scipy's STFT with the same periodic hann window, reflect padding and no scaling
    pad = np.pad(x[0].numpy(), 128, mode="reflect")
    _, _, z = scipy.signal.stft(pad, window=W.hann(256, sym=False), nperseg=256, noverlap=192, boundary=None, padded=False)
    np.testing.assert_allclose(spec[0].numpy(), np.abs(z * W.hann(256, sym=False).sum()) ** 2, rtol=1e-8, atol=1e-8)
    assert A.features.Spectrogram(256, hop_length=None)(x.float()).shape == (2, 129, 4000 // 64 + 1)   # hop = win // 4
    assert A.features.Spectrogram(256, 64, win_length=128, window=("tukey", 0.25), center=False)(x.float()).shape == (2, 129, (4000 - 256) // 64 + 1)
    mel_layer = A.features.MelSpectrogram(sr, 256, 64, n_mels=32, f_min=0.0, dtype="float64")
    mel = mel_layer(x)
    torch.testing.assert_close(mel, torch.matmul(AF.compute_fbank_matrix(sr, 256, 32, 0.0, sr // 2, dtype="float64"), spec))
    assert "fbank_matrix" not in mel_layer.state_dict()              # derived constants are not checkpointed
    logmel = A.features.LogMelSpectrogram(sr, 256, 64, n_mels=32, f_min=0.0, top_db=60.0, dtype="float64")(x)
    torch.testing.assert_close(logmel, AF.power_to_db(mel, top_db=60.0))
    mfcc = A.features.MFCC(sr, 13, 256, 64, n_mels=32, f_min=0.0, top_db=60.0, dtype="float64")(x)
    assert mfcc.shape == (2, 13, mel.shape[-1])
    np.testing.assert_allclose(mfcc.numpy(), scipy.fft.dct(logmel.numpy(), type=2, norm="ortho", axis=1)[:, :13], atol=1e-8)
    with pytest.raises(AssertionError):
        A.features.MFCC(sr, 64, n_mels=32)
    # a pure tone lands in the mel band that contains it
    t = torch.arange(sr, dtype=torch.float32) / sr
    tone = torch.sin(2 * math.pi * 1000.0 * t)[None]
    m = A.features.MelSpectrogram(sr, 512, 160, n_mels=40, f_min=0.0)(tone).mean(-1)[0]
    centres = AF.mel_frequencies(42, 0.0, sr / 2)[1:-1]
    assert abs(float(centres[int(m.argmax())]) - 1000.0) < 120.0


def test_wave_backend_and_datasets(tmp_path):
    sr = 8000
    t = torch.arange(sr) / sr
    wav = torch.stack([0.5 * torch.sin(2 * math.pi * 440 * t), 0.25 * torch.cos(2 * math.pi * 220 * t)])
    path = str(tmp_path / "a.wav")
    A.save(path, wav, sr)
    meta = A.info(path)
    assert (meta.sample_rate, meta.num_samples, meta.num_channels, meta.bits_per_sample, meta.encoding) == (sr, sr, 2, 16, "PCM_S")
    with wave.open(path) as f:
        assert f.getnchannels() == 2 and f.getsampwidth() == 2 and f.getframerate() == sr
    back, sr2 = A.load(path)
    assert sr2 == sr and back.shape == (2, sr) and back.dtype == torch.float32
    assert float((back - wav).abs().max()) <= 1.0 / 2 ** 15 + 1e-7
    raw, _ = A.load(path, normalize=False)
    torch.testing.assert_close(raw, back * 2 ** 15)
    part, _ = A.load(path, frame_offset=100, num_frames=50, channels_first=False)
    assert part.shape == (50, 2)
    torch.testing.assert_close(part, back[:, 100:150].t())
    A.save(str(tmp_path / "b.wav"), wav.t(), sr, channels_first=False)
    torch.testing.assert_close(A.load(str(tmp_path / "b.wav"))[0], back)
    with open(path, "rb") as fh:
        assert A.load(fh)[0].shape == (2, sr)
    (tmp_path / "junk.wav").write_bytes(b"not a wave file")
    with pytest.raises(NotImplementedError):
        A.load(str(tmp_path / "junk.wav"))
    assert A.backends.list_available_backends() == ["wave_backend"] and A.backends.get_current_backend() == "wave_backend"
    with pytest.raises(NotImplementedError):
        A.backends.set_backend("soundfile")
    # ESC-50 layout: meta csv + audio dir; fold `split` is the dev set
    root = tmp_path / "data"
    os.makedirs(root / "ESC-50-master" / "audio")
    os.makedirs(root / "ESC-50-master" / "meta")
    lines = ["filename,fold,target,category,esc10,src_file,take"]
    for i in range(10):
        name = f"{i % 5 + 1}-{i}-A-{i % 3}.wav"
        A.save(str(root / "ESC-50-master" / "audio" / name), wav[:1, :2000], sr)
        lines.append(f"{name},{i % 5 + 1},{i % 3},cat{i % 3},False,{i},A")
    (root / "ESC-50-master" / "meta" / "esc50.csv").write_text("\n".join(lines) + "\n")
    train = A.datasets.ESC50(mode="train", split=2, data_dir=str(root))
    dev = A.datasets.ESC50(mode="dev", split=2, data_dir=str(root))
    assert len(train) == 8 and len(dev) == 2
    feat, label = dev[0]
    assert feat.shape == (2000,) and label.dtype == np.int64 and int(label) == 1
    mf = A.datasets.ESC50(mode="dev", split=2, data_dir=str(root), feat_type="mfcc", n_mfcc=13, n_fft=256, n_mels=32)
    assert mf[0][0].shape[0] == 13
    with pytest.raises(RuntimeError):
        A.datasets.ESC50(data_dir=str(tmp_path / "nowhere"))
    with pytest.raises(RuntimeError):
        A.datasets.ESC50(data_dir=str(root), feat_type="wavelet")
    # TESS layout: <speaker>_<word>_<emotion>.wav, round-robin folds
    troot = root / "TESS_Toronto_emotional_speech_set" / "OAF_angry"
    os.makedirs(troot)
    for i, emo in enumerate(["angry", "happy", "sad", "ps", "fear", "neutral"]):
        A.save(str(troot / f"OAF_word{i}_{emo}.wav"), wav[:1, :1000], sr)
    tr = A.datasets.TESS(mode="train", n_folds=3, split=1, data_dir=str(root))
    dv = A.datasets.TESS(mode="dev", n_folds=3, split=1, data_dir=str(root))
    assert len(tr) == 4 and len(dv) == 2 and {int(dv[i][1]) for i in range(2)} <= set(range(7))
