"""Deterministic synthetic work-units and orbital template banks (numpy only, no reference data).

The reference ships one real work-unit and no expected outputs; tests and the
benchmark therefore run on synthetic inputs with the same statistics as a whitened work-unit:
zero-mean, unit-variance Gaussian noise plus a weak binary pulsar, quantised to the reference's 8-bit
``.binary`` container (``demod_binary.c:838-841``: sample = int8 / header.scale) so that exactly the
same float32 series can also be fed to the unmodified reference command line.

File formats written here follow ``structs.h:74-105`` (packed 1168-byte ``DD_Header``) and the bank
line format ``"%lg %lg %lg\\n"`` = P_orb[s] tau[s] Psi0[rad] (``demod_binary.c:197``).
"""
from __future__ import annotations

import struct
from dataclasses import dataclass

import numpy as np

TSAMPLE_US = 65.47619047619051  # sample time of the bundled PALFA work-unit
SCALE = 20.0                    # int8 counts per unit sigma
SEED = 20240001


@dataclass(frozen=True)
class Pulsar:
    """Injected binary pulsar: spin frequency, pulse duty cycle (FWHM, in turns) and circular orbit."""
    f_spin: float = 173.7
    duty: float = 0.08
    P_orb: float = 1462.994097917309
    tau: float = 0.192481315985
    Psi0: float = 1.753485476554
    summed_power: float = 120.0  # target 16-harmonic summed power (noise mean = 16) at the matching template


def _pulse_coeffs(duty: float, nharm: int = 16) -> np.ndarray:
    """|Fourier coefficients| c_m (m = 1..nharm) of a unit-height Gaussian pulse of FWHM `duty` turns,
    for the expansion pulse(phi) = c_0 + sum_m c_m cos(m phi)."""
    sigma = duty / 2.3548200450309493
    m = np.arange(1, nharm + 1)
    return 2.0 * sigma * np.sqrt(2.0 * np.pi) * np.exp(-2.0 * (np.pi * m * sigma) ** 2)


def make_series(nsamples: int, tsample_us: float = TSAMPLE_US, seed: int = SEED, pulsar: Pulsar | None = Pulsar(),
                scale: float = SCALE) -> tuple[np.ndarray, np.ndarray]:
    """Return (series float32[nsamples], int8 payload).  series == int8 / scale exactly as the reference unpacks it."""
    rng = np.random.default_rng(seed)
    x = rng.standard_normal(nsamples)
    if pulsar is not None:
        dt = tsample_us * 1e-6
        t_det = np.arange(nsamples, dtype=np.float64) * dt
        omega = 2.0 * np.pi / pulsar.P_orb
        s0 = np.sin(pulsar.Psi0)
        # detector time of pulsar time t: t_det = t - tau (sin(omega t + Psi0) - sin Psi0)   (resampler's model)
        t_psr = t_det.copy()
        for _ in range(4):
            t_psr = t_det + pulsar.tau * (np.sin(omega * t_psr + pulsar.Psi0) - s0)
        turns = pulsar.f_spin * t_psr
        frac = turns - np.floor(turns) - 0.5
        sigma = pulsar.duty / 2.3548200450309493
        profile = np.exp(-0.5 * (frac / sigma) ** 2)
        c = _pulse_coeffs(pulsar.duty)
        # harmonic m of amplitude A c_m gives normalised power A^2 c_m^2 N / 4 on top of the unit noise floor
        amp = np.sqrt(max(pulsar.summed_power - 16.0, 0.0) * 4.0 / (nsamples * np.sum(c ** 2)))
        x = x + amp * (profile - profile.mean())
    q = np.clip(np.rint(x * scale), -127, 127).astype(np.int8)
    series = (q.astype(np.float64) / scale).astype(np.float32)
    return series, q


def make_bank(n_templates: int, seed: int = SEED + 1, pulsar: Pulsar | None = Pulsar(),
              P_range=(660.0, 2231.0), tau_max: float = 0.335) -> np.ndarray:
    """Bank as float64[n,3] = (P_orb, tau, Psi0).  Row 0 is the null template `1000 0 0`, row 1 the injected
    orbit, the rest uniform draws over the range of the reference's stochastic bank.
    Values are rounded to 12 decimals so that a text round trip through the bank file is exact."""
    rng = np.random.default_rng(seed)
    bank = np.empty((n_templates, 3), np.float64)
    bank[:, 0] = rng.uniform(P_range[0], P_range[1], n_templates)
    bank[:, 1] = tau_max * np.sqrt(rng.uniform(0.0, 1.0, n_templates))
    bank[:, 2] = rng.uniform(0.0, 2.0 * np.pi, n_templates)
    if n_templates > 0:
        bank[0] = (1000.0, 0.0, 0.0)
    if n_templates > 1 and pulsar is not None:
        bank[1] = (pulsar.P_orb, pulsar.tau, pulsar.Psi0)
    return np.round(bank, 12)


def write_bank(path, bank: np.ndarray) -> None:
    with open(path, "w") as f:
        for P, tau, psi in np.asarray(bank, np.float64):
            f.write(f"{P:.12f} {tau:.12f} {psi:.12f}\n")


def read_bank(path) -> np.ndarray:
    return np.loadtxt(path, dtype=np.float64, ndmin=2)


def dd_header(nsamples: int, tsample_us: float = TSAMPLE_US, scale: float = SCALE, dm: float = 109.9) -> bytes:
    """Packed little-endian DD_Header (structs.h:74-105): 15 doubles, 3 uint32, 6 uint16, 4 x char[256]."""
    tobs = nsamples * tsample_us * 1e-6
    doubles = [tsample_us, tobs, 57310.0, 1375.0, 336.0, 61234.5, 211234.5, 187.41, -0.88, 0.0, 0.0, 0.0, 0.0, dm, scale]
    hdr = struct.pack("<15d", *doubles)
    hdr += struct.pack("<3I", nsamples, nsamples, nsamples)
    hdr += struct.pack("<6H", 1, 1, 1, 0, 0, 0)
    for text in (b"SYNTHETIC", b"synthetic.wapp", b"p0000", b"nobody"):
        hdr += text.ljust(256, b"\0")
    assert len(hdr) == 1168
    return hdr


def write_binary_wu(path, payload_int8: np.ndarray, tsample_us: float = TSAMPLE_US, scale: float = SCALE) -> None:
    """8-bit `.binary` work-unit (the path must contain ".binary": demod_binary.c:318)."""
    assert ".binary" in str(path)
    with open(path, "wb") as f:
        f.write(dd_header(len(payload_int8), tsample_us, scale))
        f.write(np.ascontiguousarray(payload_int8, np.int8).tobytes())


def write_bin4_wu(path, nibbles: np.ndarray, tsample_us: float = TSAMPLE_US, scale: float = SCALE) -> None:
    """4-bit `.bin4` work-unit: two samples per byte, high nibble first (demod_binary.c:835-836)."""
    assert ".bin4" in str(path)
    nib = np.asarray(nibbles, np.uint8)
    assert nib.size % 2 == 0 and nib.max(initial=0) < 16
    packed = (nib[0::2] << 4) | nib[1::2]
    with open(path, "wb") as f:
        f.write(dd_header(nib.size, tsample_us, scale))
        f.write(packed.tobytes())
