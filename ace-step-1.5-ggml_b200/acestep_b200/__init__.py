"""Host-side Python for the B200-native acestep_ggml hot path (ctypes bridge, sample sharding)."""
