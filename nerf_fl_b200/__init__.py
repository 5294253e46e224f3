"""B200-native NeRF-W per-ray hot path."""
