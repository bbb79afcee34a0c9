"""gravomg-b200: B200-native hierarchy construction for Gravo MG.

Submodules: capi (ctypes binding of the C ABI in include/gravomg_b200.h), sharded (row-sharded build and
solver, one rank per GPU), dump (headless .npy/.ply output), meshgen (procedural inputs), build (in-tree nvcc build).
The host-side mirror of the reference's entry points is the C++ facade under facade/."""
__version__ = "0.1.0"
