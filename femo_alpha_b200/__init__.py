"""femo_alpha_b200: B200-native (sm_100a CUDA) implementation of femo_alpha's Reissner-Mindlin
shell hot path behind the reference's Python operator surface.  See DESIGN.md."""
__version__ = "0.1.0"
