"""Audio constants of the spectral path -- same names and values as the reference's
``src/audio_parameters.py:1-20`` so that ``import audio_parameters as ap`` call sites keep working.
The CUDA library has these compiled in (include/unast_spectral.h); ``n_iter`` and ``power`` are
runtime arguments of the Griffin-Lim entry point."""
num_mels = 80
n_fft = 2048
sr = 22050
preemphasis = 0.97
frame_shift = 0.0125  # seconds
frame_length = 0.05  # seconds
hop_length = int(sr * frame_shift)  # 275 samples
win_length = int(sr * frame_length)  # 1102 samples
n_mels = 80
power = 1.2
min_level_db = -100
ref_level_db = 20
max_db = 100
ref_db = 20
n_iter = 60
outputs_per_step = 1
