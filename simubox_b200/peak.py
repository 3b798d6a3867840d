"""``gaussian_expansion`` with the reference signature (SimuBox/calculator/peak.py:11-17).

Host helper kept for API compatibility; ``Scatter.show_peak`` evaluates the same expression
on the GPU (csrc/peaks.cuh)."""
import numpy as np


def gaussian_expansion(array, amp, ctr, wid):
    return amp * np.exp(-(((array - ctr) / wid) ** 2))
