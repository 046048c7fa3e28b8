"""SURVEY 8f-3 / BASELINE configs 1 and 5 (scaled to what one-block-per-call latency allows): the
UNMODIFIED reference encoder application (oracle/_ref/SvtAv1EncApp, built from the reference's
own sources by `make -C oracle encapp`) encodes the same synthetic clip twice -- once as it is
(C/AVX2 transform + quantiser path), once with integration/svt_gpu_preload.c rebinding its RTCD
slots and entry points to libsvtgpu.so -- and the two .ivf bitstreams must be byte-identical.

Test infrastructure: the encoder binaries live under oracle/_ref/ (git-ignored, they travel to the
GPU box with the snapshot); /root/reference is not needed at run time."""
import hashlib
import os
import subprocess
import time

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REFDIR = os.path.join(ROOT, "oracle", "_ref")
APP = os.path.join(REFDIR, "SvtAv1EncApp")
PRELOAD = os.path.join(REFDIR, "libsvtgpu_preload.so")
GPULIB = os.path.join(ROOT, "cidana_svt_av1_b200", "libsvtgpu.so")


def make_clip(path, w, h, n, seed=5, bd=8):
    """Moving sinusoid pattern + noise, 4:2:0 planar; 8-bit bytes, or 10-bit samples in
    little-endian 16-bit words (the application's unpacked 10-bit input format)."""
    rng = np.random.default_rng(seed)
    yy, xx = np.mgrid[0:h, 0:w]
    k = 1 << (bd - 8)
    with open(path, "wb") as f:
        for t in range(n):
            y = 128 + 60 * np.sin((xx + 3 * t) / 17.0) + 40 * np.cos((yy - 2 * t) / 11.0) + rng.normal(0, 6, (h, w))
            u = 128 + 30 * np.sin((xx[::2, ::2] + t) / 23.0)
            v = 128 + 30 * np.cos((yy[::2, ::2] + t) / 19.0)
            for p in (y, u, v):
                f.write(np.clip(p * k, 0, 256 * k - 1).astype(np.uint8 if bd == 8 else "<u2").tobytes())


def encode(clip, out, w, h, n, qp, gpu_level=None, timeout=900, bd=8):
    env = dict(os.environ)
    if gpu_level:
        env.update(LD_PRELOAD=PRELOAD, SVT_GPU_LIB=GPULIB, SVT_GPU_LEVEL=gpu_level)
    cmd = [APP, "-i", clip, "-w", str(w), "-h", str(h), "-n", str(n), "-enc-mode", "8", "-q", str(qp),
           "-b", out] + (["-bit-depth", str(bd)] if bd != 8 else [])
    t0 = time.time()
    r = subprocess.run(cmd, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True,
                       timeout=timeout)
    assert r.returncode == 0 and os.path.getsize(out) > 32, r.stdout[-2000:]
    return time.time() - t0, r.stdout


def md5(path):
    return hashlib.md5(open(path, "rb").read()).hexdigest()


needs_app = pytest.mark.skipif(not (os.path.exists(APP) and os.path.exists(PRELOAD)),
                               reason="oracle/_ref/SvtAv1EncApp not built (make -C oracle encapp needs /root/reference)")


@needs_app
def test_reference_encoder_is_deterministic(tmp_path):
    """Config 1 at small scale, CPU only: the oracle application itself, twice."""
    clip = str(tmp_path / "clip.yuv")
    make_clip(clip, 192, 128, 3)
    encode(clip, str(tmp_path / "a.ivf"), 192, 128, 3, 30)
    encode(clip, str(tmp_path / "b.ivf"), 192, 128, 3, 30)
    assert md5(str(tmp_path / "a.ivf")) == md5(str(tmp_path / "b.ivf"))


@needs_app
def test_preload_fails_loudly_without_the_library(tmp_path):
    clip = str(tmp_path / "clip.yuv")
    make_clip(clip, 192, 128, 2)
    env = dict(os.environ, LD_PRELOAD=PRELOAD, SVT_GPU_LIB=str(tmp_path / "missing.so"))
    r = subprocess.run([APP, "-i", clip, "-w", "192", "-h", "128", "-n", "2", "-enc-mode", "8", "-b",
                        str(tmp_path / "x.ivf")], env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT,
                       text=True, timeout=120)
    assert r.returncode != 0 and "cannot load libsvtgpu.so" in r.stdout


@needs_app
def test_reference_encoder_is_deterministic_10bit(tmp_path):
    clip = str(tmp_path / "clip.yuv")
    make_clip(clip, 192, 128, 3, bd=10)
    encode(clip, str(tmp_path / "a.ivf"), 192, 128, 3, 30, bd=10)
    encode(clip, str(tmp_path / "b.ivf"), 192, 128, 3, 30, bd=10)
    assert md5(str(tmp_path / "a.ivf")) == md5(str(tmp_path / "b.ivf"))


@pytest.mark.gpu
@needs_app
@pytest.mark.parametrize("level,w,h,n,qp,bd", [("A", 320, 192, 6, 30, 8), ("AB", 320, 192, 6, 30, 8),
                                               ("AB", 640, 384, 4, 45, 8),
                                               # S: encode_pass_tx_search batched, one launch per block
                                               ("ABS", 320, 192, 6, 30, 8), ("ABS", 640, 384, 4, 45, 8),
                                               ("BS", 320, 192, 4, 20, 8),
                                               # 10-bit: Av1EncodeLoop16bit (EbCodingLoop.c:1020) ->
                                               # av1_inv_transform_recon (EbTransforms.c:8905), and mode
                                               # decision quantising chroma with the 8-bit *Md tables
                                               ("A", 320, 192, 4, 30, 10), ("B", 320, 192, 4, 30, 10),
                                               ("AB", 320, 192, 4, 45, 10)])
def test_bitstream_identical_with_gpu_path(tmp_path, level, w, h, n, qp, bd):
    clip = str(tmp_path / "clip.yuv")
    make_clip(clip, w, h, n, bd=bd)
    ref, gpu = str(tmp_path / "ref.ivf"), str(tmp_path / "gpu.ivf")
    t_ref, _ = encode(clip, ref, w, h, n, qp, bd=bd)
    t_gpu, log = encode(clip, gpu, w, h, n, qp, gpu_level=level, bd=bd)
    assert "RTCD slots rebound" in log, log[-2000:]
    print(f"\n{w}x{h} x{n} q{qp} {bd}-bit level {level}: reference {t_ref:.1f} s, GPU path {t_gpu:.1f} s, "
          f"{os.path.getsize(ref)} bytes\n" + "\n".join(l for l in log.splitlines() if "svtgpu" in l))
    assert open(ref, "rb").read() == open(gpu, "rb").read()
    if "S" in level:  # the batched search really ran, with more than one candidate per block
        served = [l for l in log.splitlines() if "batched tx-type searches" in l]
        n_search = int(served[-1].split("batched tx-type searches ")[1].split(" ")[0])
        n_cand = int(served[-1].split("(")[-1].split(" ")[0])
        assert n_search > 0 and n_cand > 2 * n_search, served
    if bd == 10 and "B" in level:  # the 16-bit reconstruction entry point was really driven
        served = [l for l in log.splitlines() if "level-B calls served" in l]
        assert served and int(served[-1].split("inv_transform_recon ")[1].split(",")[0]) > 0, served
