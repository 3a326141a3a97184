"""B200-native optimisation kernel for IGM (A/M hot path)."""
