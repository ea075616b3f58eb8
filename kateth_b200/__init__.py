"""kateth_b200 -- host-side mirror of kateth's `Setup<4096, 65>` over the C ABI of the B200-native KZG engine.

The reference is Rust (no toolchain in this image); this thin ctypes layer exists so the parity tests read like
kateth's own (/root/reference/src/kzg/setup.rs:319-463): same method names, argument order and error behaviour.
All computation happens in libkateth_b200.so (hand-written sm_100a CUDA + a host pairing); there is NO CPU
fallback -- without the built library or without a GPU, loading a Setup raises.
"""
from .setup import Setup, Group, KzgError, BatchVerifier, batch_error_of_records, shard_bounds, set_table_cache, g1_validate, lib, lib_path, BYTES_PER_BLOB, RECORD_BYTES, test_hooks  # noqa: F401
