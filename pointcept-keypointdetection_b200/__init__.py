"""ptv3_b200: B200-native (sm_100a) PTv3 backbone hot path behind the reference's operator surface."""
__version__ = "0.1.0"
