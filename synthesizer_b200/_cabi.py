"""
ctypes binding of libsynth_b200.so (declarations: include/synth_b200.h).

This is the only way the Python host code reaches the GPU: there is no CPU fallback and no
alternative backend.  If the library is missing, importing this module raises — build it
with ``python -c "import __graft_entry__ as g; g.build()"`` or synthesizer_b200/csrc/build.sh.
"""
import ctypes
import os
from ctypes import (POINTER, Structure, c_char_p, c_double, c_float, c_int, c_int32, c_int64,
                    c_uint, c_uint64, c_ulonglong, c_void_p)

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("SYNTH_B200_LIB") or os.path.join(_HERE, "libsynth_b200.so")

WAVE_NONE, WAVE_SINE, WAVE_TRIANGLE, WAVE_SQUARE, WAVE_SAWTOOTH, WAVE_PULSE = -1, 0, 1, 2, 3, 4
WAVE_SEMICIRCLE, WAVE_POINTY = 5, 6
TIMEBASE_MAX_SEGS = 256
ECHO_MAX = 64


class SbError(RuntimeError):
    pass


class TimeSeg(Structure):
    _fields_ = [("n_start", c_int64), ("t_start", c_double), ("step", c_double)]


class Envelope(Structure):
    _fields_ = [("n_a", c_int64), ("n_d", c_int64), ("n_s", c_int64), ("n_r", c_int64),
                ("extra", c_int32), ("enabled", c_int32),
                ("a_slope", c_double), ("d_slope", c_double), ("r_slope", c_double),
                ("sustain_level", c_double), ("amp_extra", c_double)]


class OscParams(Structure):
    _fields_ = [("wave", c_int32), ("reserved", c_int32),
                ("frequency", c_double), ("amplitude", c_double), ("phase", c_double),
                ("bias", c_double), ("pulsewidth", c_double), ("samplerate", c_double)]


class Voice(Structure):
    _fields_ = [("c_wave", c_int32), ("m_wave", c_int32),
                ("c_freq", c_double), ("c_phase", c_double),
                ("m_freq", c_double), ("m_phase", c_double),
                ("m_amp", c_double), ("m_bias", c_double),
                ("c_pw", c_double), ("m_pw", c_double),
                ("c_amp", c_float), ("c_bias", c_float),
                ("gain_l", c_float), ("gain_r", c_float),
                ("env", Envelope)]


class BankPlan(Structure):
    _fields_ = [("nvoices", c_int32), ("voices_per_group", c_int32), ("ngroups", c_int32),
                ("nspans", c_int32), ("frames", c_int64), ("span_len", c_int64),
                ("fast_span_len", c_int64), ("fast_nspans", c_int32), ("nsegs_rad", c_int32),
                ("generic_voices", c_int32), ("nprefix", c_int32), ("edged_voices", c_int32),
                ("scan_voices", c_int32), ("n0", c_int64),
                ("off_fastv", c_int64), ("off_tab", c_int64),
                ("carry_bytes", c_int64), ("partial_bytes", c_int64), ("smooth_voices", c_int32)]


if not os.path.exists(LIB_PATH):
    raise ImportError(
        "libsynth_b200.so is not built (%s). There is no CPU fallback: run "
        "synthesizer_b200/csrc/build.sh (needs nvcc; cross-compiles sm_100a without a GPU)." % LIB_PATH)

lib = ctypes.CDLL(LIB_PATH)

_P = c_void_p
_SIGNATURES = {
    # name: (restype, argtypes)
    "sb_abi_version": (c_int, []),
    "sb_last_error": (c_char_p, []),
    "sb_launch_count": (c_int64, []),
    "sb_timebase_build": (c_int, [c_double, c_int64, POINTER(TimeSeg), c_int]),
    "sb_timebase_build_from": (c_int, [c_double, c_double, c_int64, POINTER(TimeSeg), c_int]),
    "sb_harmonics_render_f32": (c_int, [POINTER(OscParams), POINTER(c_int32), POINTER(c_float), c_int, _P, c_int,
                                        c_int64, c_int64, _P, _P, _P]),
    "sb_timebase_at": (c_double, [POINTER(TimeSeg), c_int, c_int64]),
    "sb_timebase_first_ge": (c_int64, [POINTER(TimeSeg), c_int, c_int64, c_double]),
    "sb_envelope_build": (c_int, [c_double, c_double, c_double, c_double, c_double, c_double, POINTER(Envelope)]),
    "sb_osc_render_f32": (c_int, [POINTER(OscParams), _P, c_int, c_int64, c_int64, _P, _P, _P, _P]),
    "sb_fm_prefix_scratch_bytes": (c_int64, [c_int64]),
    "sb_fm_prefix_f64": (c_int, [_P, _P, c_int, c_int64, c_int64, _P, _P, _P, _P]),
    "sb_fm_prefix_f64_from_f64": (c_int, [_P, _P, c_int, c_int64, c_int64, _P, _P, _P, _P]),
    "sb_osc_render_f64": (c_int, [POINTER(OscParams), _P, c_int, c_int64, c_int64, _P, _P, _P, _P]),
    "sb_fm_prefix_noalloc_max_count": (c_int64, []),
    "sb_fm_prefix_f64_noalloc": (c_int, [_P, _P, c_int, c_int64, c_int64, _P, _P, _P]),
    "sb_noise_render_f32": (c_int, [_P, c_int64, c_int64, c_double, c_double, c_int64, c_int64, _P, _P]),
    "sb_linear_render_f32": (c_int, [c_double, c_double, c_double, c_double, c_int64, c_int64, _P, _P]),
    "sb_linear_render_f64": (c_int, [c_double, c_double, c_double, c_double, c_int64, c_int64, _P, _P]),
    "sb_uniform_f32": (c_int, [c_uint64, c_int64, c_int64, _P, _P]),
    "sb_envelope_apply_f32": (c_int, [_P, POINTER(Envelope), c_int64, c_int64, _P, _P]),
    "sb_envelope_apply_f64": (c_int, [_P, POINTER(Envelope), c_int64, c_int64, _P, _P]),
    "sb_mul_f32": (c_int, [_P, _P, c_int64, _P, _P]),
    "sb_mix_f32": (c_int, [_P, c_int, c_int64, _P, _P]),
    "sb_clip_f32": (c_int, [_P, c_float, c_float, c_int64, _P, _P]),
    "sb_abs_f32": (c_int, [_P, c_int64, _P, _P]),
    "sb_scale_bias_f32": (c_int, [_P, c_float, c_float, c_int64, _P, _P]),
    "sb_channel_f32": (c_int, [_P, c_int, c_int, c_int64, _P, _P]),
    "sb_echo_f32": (c_int, [_P, c_int64, c_int64, c_int, POINTER(c_int64), POINTER(c_float), c_int64, c_int64, _P, _P]),
    "sb_f32_to_pcm16": (c_int, [_P, c_int64, c_float, _P, _P, _P]),
    "sb_f32_to_pcm32": (c_int, [_P, c_int64, c_double, _P, _P, _P]),
    "sb_f32x2_to_pcm16_stereo": (c_int, [_P, _P, c_int64, c_float, _P, _P, _P]),
    "sb_bank_plan_make": (c_int, [c_int, c_int64, c_int64, c_int, c_int, POINTER(BankPlan)]),
    "sb_bank_plan_hint": (c_int, [POINTER(BankPlan), c_int, c_int, c_int, c_int]),
    "sb_bank_count_classes": (c_int, [_P, c_int, c_double, POINTER(c_int)]),
    "sb_voice_is_fast": (c_int, [POINTER(Voice), c_double]),
    "sb_bank_render_f32": (c_int, [_P, POINTER(BankPlan), _P, c_int, _P, c_int, c_int64, c_double, _P, _P, _P, _P]),
    "sb_pcm16_add": (c_int, [_P, _P, c_int64, _P, _P]),
    "sb_pcm16_mul": (c_int, [_P, c_int64, c_double, _P, _P]),
    "sb_pcm16_bias": (c_int, [_P, c_int64, c_int, _P, _P]),
    "sb_pcm16_reverse": (c_int, [_P, c_int64, _P, _P]),
    "sb_pcm16_maxabs": (c_int, [_P, c_int64, _P, _P]),
    "sb_pcm16_tomono": (c_int, [_P, c_int64, c_double, c_double, _P, _P]),
    "sb_pcm16_tostereo": (c_int, [_P, c_int64, c_double, c_double, _P, _P]),
    "sb_pcm16_to_pcm32": (c_int, [_P, c_int64, _P, _P]),
    "sb_pcm32_to_pcm16": (c_int, [_P, c_int64, _P, _P]),
    "sb_pcm16_fadeout": (c_int, [_P, c_int64, c_double, _P]),
    "sb_pcm16_fadein": (c_int, [_P, c_int64, c_double, _P]),
    "sb_bank_render_check": (c_int, [_P, _P, _P, _P]),
    "sb_pcm16_modulate": (c_int, [_P, _P, c_int64, _P]),
    "sb_pcm16_modulate_cycle_f64": (c_int, [_P, _P, c_int64, c_int64, _P]),
    "sb_pcm32_modulate_cycle_f64": (c_int, [_P, _P, c_int64, c_int64, _P]),
    "sb_pcm16_ratecv_frames": (c_int64, [c_int64, c_int64, c_int64]),
    "sb_pcm16_ratecv": (c_int, [_P, c_int64, c_int, c_int64, c_int64, _P, _P]),
    "sb_pcm16_ratecv_many": (c_int, [_P, _P, _P, _P, _P, _P, _P, c_int64, c_int, c_int64, _P, _P]),
    "sb_pcm16_mix_at_many_scratch_bytes": (c_int64, [c_int64, c_int64, c_int64]),
    "sb_pcm16_mix_at_many": (c_int, [_P, c_int64, _P, _P, _P, _P, c_int64, _P, c_int64, _P]),
    "sb_pcm32_add": (c_int, [_P, _P, c_int64, _P, _P]),
    "sb_pcm32_mul": (c_int, [_P, c_int64, c_double, _P, _P]),
    "sb_pcm32_bias": (c_int, [_P, c_int64, c_int, _P, _P]),
    "sb_pcm32_reverse": (c_int, [_P, c_int64, _P, _P]),
    "sb_pcm32_maxabs": (c_int, [_P, c_int64, _P, _P]),
    "sb_pcm32_tomono": (c_int, [_P, c_int64, c_double, c_double, _P, _P]),
    "sb_pcm32_tostereo": (c_int, [_P, c_int64, c_double, c_double, _P, _P]),
    "sb_pcm32_fadeout": (c_int, [_P, c_int64, c_double, _P]),
    "sb_pcm32_fadein": (c_int, [_P, c_int64, c_double, _P]),
    "sb_pcm32_modulate": (c_int, [_P, _P, c_int64, _P]),
    "sb_pcm32_ratecv": (c_int, [_P, c_int64, c_int, c_int64, c_int64, _P, _P]),
    "sb_pcm24_unpack": (c_int, [_P, c_int64, _P, _P]),
    "sb_pcm24_pack": (c_int, [_P, c_int64, _P, _P]),
    "sb_pcm24_clamp": (c_int, [_P, c_int64, _P]),
    "sb_pcm32_shift": (c_int, [_P, c_int64, c_int, _P]),
    "sb_host_bank_render_pcm16": (c_int, [POINTER(Voice), c_int, c_int64, c_double, c_float, _P, POINTER(c_ulonglong)]),
    "sb_host_pcm16_mix_at_many": (c_int, [_P, c_int64, _P, c_int64, _P, _P, _P, c_int64]),
    "sb_host_bank_render_pcm16_on_stream": (c_int, [POINTER(Voice), c_int, c_int64, c_double, c_float, _P,
                                                    POINTER(c_ulonglong), _P]),
    "sb_host_bank_render_mix_f32": (c_int, [POINTER(Voice), c_int, c_int64, c_double, _P, _P]),
    "sb_host_reserve": (c_int, [c_int, c_int64, c_double]),
    "sb_host_pcm16_mix_at_many_on_stream": (c_int, [_P, c_int64, _P, c_int64, _P, _P, _P, c_int64, _P]),
    "sb_bank_profile_events": (c_int, [_P, _P]),
    "sb_event_create": (c_int, [POINTER(c_void_p)]),
    "sb_event_elapsed_ms": (c_int, [_P, _P, POINTER(c_float)]),
    "sb_event_destroy": (c_int, [_P]),
}

EXPORTED = tuple(_SIGNATURES)

for _name, (_res, _args) in _SIGNATURES.items():
    _fn = getattr(lib, _name)          # AttributeError here = header and library disagree
    _fn.restype = _res
    _fn.argtypes = _args


def check(rc, what=""):
    """Raise SbError for a negative return code, with the library's message."""
    if rc < 0:
        msg = lib.sb_last_error()
        raise SbError("%s failed (%d): %s" % (what or "libsynth_b200 call", rc,
                                              msg.decode("utf-8", "replace") if msg else "?"))
    return rc


def call(name, *args):
    return check(getattr(lib, name)(*args), name)
