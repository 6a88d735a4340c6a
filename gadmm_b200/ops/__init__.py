"""ctypes bindings of the native sm_100a library (``libgadmm_b200.so``, C ABI in
``include/gadmm_b200.h``): :mod:`native` (loader, descriptor structs), :mod:`linalg` (Gram,
SPD inverse, packing, streaming SYMV), :mod:`topology` (device findPath, Philox fills)."""
