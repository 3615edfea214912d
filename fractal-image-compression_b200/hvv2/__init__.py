"""HV-partition fractal codec, B200 back end.  Mirrors the reference package `hvv2`."""
