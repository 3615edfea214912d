"""Quadtree fractal codec, B200 back end.  Mirrors the reference package `quadtree`."""
