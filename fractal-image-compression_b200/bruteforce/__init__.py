"""Brute-force fractal codec, B200 back end.  Mirrors the reference package `bruteforce`
(bruteforce/bf_encoder.py, jit_bfencoder.py, decoder.py, model.py)."""
