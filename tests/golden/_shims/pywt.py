"""Fixture: the four 2-tap Haar filter lists the reference reads from PyWavelets
(DWT_IDWT/DWT_IDWT_layer.py:262-264, 352-356).  PyWavelets is not installed here."""
import math


class Wavelet:
    def __init__(self, name):
        if name != "haar":
            raise ValueError("fixture only provides the haar wavelet")
        s = 1.0 / math.sqrt(2.0)
        self.dec_lo, self.dec_hi = [s, s], [-s, s]
        self.rec_lo, self.rec_hi = [s, s], [s, -s]
