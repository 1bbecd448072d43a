"""B200-native RNS-CKKS engine for the encrypted sorts of oksuman/FHE-Sorting (hot path only)."""
from .lib import Context, Ciphertext, Plaintext, Keys, FheError, load, params_chain, launch_count, algo_bytes, LIB_PATH  # noqa: F401
