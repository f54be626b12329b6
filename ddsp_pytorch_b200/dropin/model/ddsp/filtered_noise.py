from ddsp_pytorch_b200.modules import FilteredNoise, amp_to_impulse_response, fft_convolve  # noqa: F401
