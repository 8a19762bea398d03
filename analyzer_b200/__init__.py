"""analyzer_b200 -- B200-native TrueSkill rating hot path of vainglorygame/analyzer.

    analyzer_b200.rater    drop-in for the reference's rater module
    analyzer_b200.engine   env / HBM player table / schedule / rate over the C ABI
    analyzer_b200._lib     ctypes binding of libtsrate.so (include/tsrate.h)
"""
__all__ = ["rater", "engine"]
