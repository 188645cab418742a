"""B200-native batched RWKV-7 forward step (drop-in for Albatross rwkv7.py behind Chirrup)."""
