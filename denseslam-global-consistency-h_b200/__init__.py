"""B200-native volumetric-fusion hot path behind InfiniTAM's engine interfaces.

Layout: csrc/ (sm_100a CUDA kernels + the C ABI of include/dsb200.h -> libdsb200.so),
abi.py (ctypes mirror of the C ABI), engine.py (host-side mirror of the reference's
ITMSceneReconstructionEngine / ITMVisualisationEngine call surface), synth.py (deterministic
synthetic depth / pose generator).  Import with importlib (the directory name has hyphens):
    pkg = importlib.import_module("denseslam-global-consistency-h_b200")
"""
