"""kair_b200: B200 (sm_100a) kernels for KAIR's image-restoration hot path (SwinIR / DnCNN / DRUNet)."""
__version__ = "0.1.0"
