"""first_raytracer_b200 — B200-native ray-tracing core for jammm/first_raytracer's path integrator.

csrc/  hand-written sm_100a CUDA (wavefront generate/extend/shade/connect + ray casting) → libfrt.so
host/  C++ host layer with the reference's Scene/hitable/camera/integrator API → libfrt_host.so, frt_main
api.py ctypes binding used by tests/ and bench.py
"""
__all__ = ["api"]
