"""unast_b200 -- B200-native (sm_100a) spectral signal path of Lucaskabela/UNAST.

Public surface mirrors the reference modules this path replaces:
    unast_b200.utils            get_spectrograms / spectrogram2wav / griffin_lim / invert_spectrogram
    unast_b200.gl_vocoder       vocode(list_file, mels_dir, out_dir)
    unast_b200.prepare_data     PrepareDataset / prepare
    unast_b200.audio_parameters the constants module (``ap``)
    unast_b200.engine           batched device-side API over libunast_spectral.so

Importing the package does not touch the GPU; the first call that needs it loads the shared library
and fails loudly if it (or a CUDA device) is missing.
"""
__version__ = "0.1.0"
