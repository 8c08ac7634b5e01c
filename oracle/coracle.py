"""
ORACLE — TEST INFRASTRUCTURE ONLY.  ctypes loader for oracle/_build/liboracle.so (the C
restatement in oracle.c).  Builds it with the Makefile when missing.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "liboracle.so")

SINE, TRIANGLE, SQUARE, SAWTOOTH, PULSE = range(5)
VOICE_STRIDE = 18


def build(force=False):
    src = os.path.join(_HERE, "oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-s", "-C", _HERE, "-B" if force else "-s"])
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = ctypes.CDLL(_SO)
        L = _lib
        d, ll, i, p = ctypes.c_double, ctypes.c_longlong, ctypes.c_int, ctypes.c_void_p
        L.orc_osc.argtypes = [i, ll, d, d, d, d, d, p, d, p, p]
        L.orc_osc.restype = None
        L.orc_noise.argtypes = [ll, d, d, d, d, p, p]
        L.orc_envelope.argtypes = [p, ll, d, d, d, d, d, d, i, p]
        L.orc_envelope.restype = ll
        L.orc_echo.argtypes = [p, ll, d, d, i, d, d, p]
        L.orc_to_pcm16.argtypes = [p, ll, d, p]
        L.orc_to_pcm16.restype = ll
        L.orc_voice_bank.argtypes = [p, ll, ll, d, p, p]
        L.orc_pcm16_add.argtypes = [p, p, ll, p]
        L.orc_pcm16_mul.argtypes = [p, ll, d, p]
        L.orc_pcm16_mix_at_many.argtypes = [p, ll, p, p, p, p, ll]
        L.orc_pcm16_ratecv.argtypes = [p, ll, i, ll, ll, p, ll]
        L.orc_pcm16_ratecv.restype = ll
        L.orc_pcm16_fadeout.argtypes = [p, ll, d]
        L.orc_pcm16_fadein.argtypes = [p, ll, d]
    return _lib


def _ptr(a):
    return a.ctypes.data_as(ctypes.c_void_p) if a is not None else None


def osc(wave, n, samplerate, frequency, amplitude=1.0, phase=0.0, bias=0.0, fm=None,
        pulsewidth=0.1, pwm=None):
    out = np.empty(n, dtype=np.float64)
    fm = None if fm is None else np.ascontiguousarray(fm, dtype=np.float64)
    pwm = None if pwm is None else np.ascontiguousarray(pwm, dtype=np.float64)
    lib().orc_osc(wave, n, samplerate, frequency, amplitude, phase, bias, _ptr(fm), pulsewidth,
                  _ptr(pwm), _ptr(out))
    return out


def noise(n, samplerate, frequency, amplitude, bias, u):
    u = np.ascontiguousarray(u, dtype=np.float64)
    out = np.empty(n, dtype=np.float64)
    lib().orc_noise(n, samplerate, frequency, amplitude, bias, _ptr(u), _ptr(out))
    return out


def envelope(x, samplerate, attack, decay, sustain, sustain_level, release, stop_at_end=False):
    x = np.ascontiguousarray(x, dtype=np.float64)
    out = np.empty_like(x)
    n = lib().orc_envelope(_ptr(x), len(x), samplerate, attack, decay, sustain, sustain_level,
                           release, int(stop_at_end), _ptr(out))
    return out[:n]


def echo(x, samplerate, after, amount, delay, decay):
    x = np.ascontiguousarray(x, dtype=np.float64)
    out = np.empty_like(x)
    lib().orc_echo(_ptr(x), len(x), samplerate, after, amount, delay, decay, _ptr(out))
    return out


def to_pcm16(x, scale=32767.0):
    x = np.ascontiguousarray(x, dtype=np.float64)
    out = np.empty(len(x), dtype=np.int16)
    bad = lib().orc_to_pcm16(_ptr(x), len(x), scale, _ptr(out))
    return out, bad


def voice_bank(params, nframes, samplerate):
    params = np.ascontiguousarray(params, dtype=np.float64).reshape(-1, VOICE_STRIDE)
    left = np.empty(nframes, dtype=np.float64)
    right = np.empty(nframes, dtype=np.float64)
    lib().orc_voice_bank(_ptr(params), params.shape[0], nframes, samplerate, _ptr(left), _ptr(right))
    return left, right


def pcm16_add(a, b):
    a = np.ascontiguousarray(a, dtype=np.int16)
    b = np.ascontiguousarray(b, dtype=np.int16)
    out = np.empty_like(a)
    lib().orc_pcm16_add(_ptr(a), _ptr(b), len(a), _ptr(out))
    return out


def pcm16_mul(a, factor):
    a = np.ascontiguousarray(a, dtype=np.int16)
    out = np.empty_like(a)
    lib().orc_pcm16_mul(_ptr(a), len(a), factor, _ptr(out))
    return out


def pcm16_mix_at_many(dst, src_pool, src_off, src_len, dst_off):
    dst = np.array(dst, dtype=np.int16, copy=True)
    src_pool = np.ascontiguousarray(src_pool, dtype=np.int16)
    so = np.ascontiguousarray(src_off, dtype=np.int64)
    sl = np.ascontiguousarray(src_len, dtype=np.int64)
    do = np.ascontiguousarray(dst_off, dtype=np.int64)
    lib().orc_pcm16_mix_at_many(_ptr(dst), len(dst), _ptr(src_pool), _ptr(so), _ptr(sl), _ptr(do), len(so))
    return dst


def pcm16_ratecv(a, nchannels, inrate, outrate):
    a = np.ascontiguousarray(a, dtype=np.int16)
    nframes = len(a) // nchannels
    cap = (nframes * outrate) // inrate + 8
    out = np.empty(cap * nchannels, dtype=np.int16)
    n = lib().orc_pcm16_ratecv(_ptr(a), nframes, nchannels, inrate, outrate, _ptr(out), cap)
    assert n >= 0
    return out[: n * nchannels]


def pcm16_fadeout(a, target_volume=0.0):
    a = np.array(a, dtype=np.int16, copy=True)
    lib().orc_pcm16_fadeout(_ptr(a), len(a), target_volume)
    return a


def pcm16_fadein(a, start_volume=0.0):
    a = np.array(a, dtype=np.int16, copy=True)
    lib().orc_pcm16_fadein(_ptr(a), len(a), start_volume)
    return a
