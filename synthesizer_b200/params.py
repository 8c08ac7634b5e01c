"""
Global defaults, mirroring synthplayer.params (recalled; source absent offline):
sample rate / width / channels that Sample.normalize() converts to, and the oscillator
block size that Oscillator.blocks() yields.
"""
norm_samplerate = 44100
norm_nchannels = 2
norm_samplewidth = 2
norm_frames_per_chunk = norm_samplerate // 30
norm_osc_blocksize = 512

# backend knobs (no upstream counterpart): frames rendered on the GPU per host round trip
# when an oscillator is consumed through blocks()/generator()
render_chunk_min = 1 << 14
render_chunk_max = 1 << 20
