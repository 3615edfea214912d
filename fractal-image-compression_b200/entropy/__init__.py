"""Entropy coding of the serialised transform fields.  Mirrors the reference package `entropy`."""
