"""gravomg-b200: B200-native hierarchy construction for Gravo MG.

Submodules: capi (ctypes binding of the C ABI in include/gravomg_b200.h), hierarchy (host-side
mirror of the reference entry points), meshgen (procedural inputs)."""
__version__ = "0.1.0"
