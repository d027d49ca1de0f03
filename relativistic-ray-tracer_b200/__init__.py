"""relativistic-ray-tracer_b200 — B200-native hot path of the relativistic ray tracer.

The product is the C-ABI library csrc/librrt_b200.so (include/rrt_b200.h).  This package is the Python-side
plumbing around it: scene containers (scene.py), the ctypes binding and the PathTracer/BVHAccel mirror of the
reference interface (api.py), the procedural scene generator (synth.py).  Import with
importlib.import_module("relativistic-ray-tracer_b200").
"""
__version__ = "0.1.0"
