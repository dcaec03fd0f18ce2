"""B200-native shallow-water right-hand-side path for RDycore (hot path only)."""
