"""CPU tier: the warp-level arithmetic of the K1 spectrogram kernel (csrc/spectrogram_core.h), emulated lane by lane,
against numpy's float64 FFT -- validates the radix-8 butterflies, twiddles, both shared-memory exchange index maps and the
in-register split of the packed frame pair, and counts shared-memory bank conflicts, with no GPU."""
import ctypes as C
import subprocess
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent
SRC = ROOT / "tests" / "emu" / "spectrogram_emu.cpp"
SO = ROOT / "tests" / "emu" / "libspectrogram_emu.so"


@pytest.fixture(scope="module")
def emu():
    core = ROOT / "nntoolkitcore_b200" / "csrc" / "spectrogram_core.h"
    if not SO.exists() or SO.stat().st_mtime < max(SRC.stat().st_mtime, core.stat().st_mtime):
        subprocess.run(["g++", "-O1", "-std=c++17", "-fPIC", "-shared", "-ffp-contract=off", "-o", str(SO), str(SRC)],
                       check=True)
    lib = C.CDLL(str(SO))
    lib.emu_spectrogram512_pair.restype = C.c_int
    return lib


def _tw():
    i = np.arange(512)
    ph = -2.0 * np.pi * i / 512
    return np.stack([np.cos(ph), np.sin(ph)], axis=1).astype(np.float32)


def _run(emu, a, b):
    tw = _tw()
    pa = np.zeros(257, np.float32); pb = np.zeros(257, np.float32); z = np.zeros((512, 2), np.float32)
    f = lambda x: x.ctypes.data_as(C.POINTER(C.c_float))
    extra = emu.emu_spectrogram512_pair(f(a), f(b), f(tw), f(pa), f(pb), f(z))
    return pa, pb, z[:, 0] + 1j * z[:, 1], extra


@pytest.mark.parametrize("seed", [0, 1, 2])
def test_emulated_warp_matches_fft(emu, seed):
    rng = np.random.default_rng(seed)
    a = rng.standard_normal(512).astype(np.float32)
    b = rng.standard_normal(512).astype(np.float32)
    pa, pb, z, _ = _run(emu, a, b)
    want_z = np.fft.fft(a.astype(np.float64) + 1j * b.astype(np.float64))
    assert np.abs(z - want_z).max() <= 2e-5 * np.abs(want_z).max()
    A = np.abs(np.fft.rfft(a.astype(np.float64))) ** 2 * 4
    B = np.abs(np.fft.rfft(b.astype(np.float64))) ** 2 * 4
    assert np.abs(pa - A).max() <= 1e-5 * A.max()
    assert np.abs(pb - B).max() <= 1e-5 * B.max()


def test_impulse_and_dc(emu):
    a = np.zeros(512, np.float32); a[37] = 1.0      # flat spectrum
    b = np.ones(512, np.float32)                     # DC only
    pa, pb, _, _ = _run(emu, a, b)
    assert np.allclose(pa, 4.0, atol=1e-4)
    assert abs(pb[0] - 4 * 512 * 512) <= 1e-6 * 4 * 512 * 512 and np.abs(pb[1:]).max() <= 1e-3


def test_exchanges_bank_conflicts(emu):
    """8+8 (exchange 1) and 8+16 (exchange 2) warp-wide instructions, all conflict-free: 160 wavefronts per frame pair."""
    rng = np.random.default_rng(5)
    a = rng.standard_normal(512).astype(np.float32)
    _, _, _, extra = _run(emu, a, a)
    assert emu.emu_spectrogram512_smem_instructions() == 40
    assert extra == 0, f"{extra} extra shared-memory wavefronts per frame pair"
