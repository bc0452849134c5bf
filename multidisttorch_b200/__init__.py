"""B200-native hot path for MultiDistTorch-style multi-group DDP (VAE-HPO trial step)."""
__version__ = "0.1.0"
