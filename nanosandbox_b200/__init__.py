"""nanosandbox_b200: B200-native nanoGPT data-parallel training step."""
__version__ = "0.1.0"
