"""fod_b200 — B200-native hot path of atonderski/future-object-detection."""
__version__ = "0.1.0"
