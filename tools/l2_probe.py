import ctypes
rt = ctypes.CDLL("libcudart.so.12")
v = ctypes.c_int()
for name, a in (("MaxPersistingL2CacheSize", 108), ("MaxAccessPolicyWindowSize", 109), ("L2CacheSize", 38)):
    rt.cudaDeviceGetAttribute(ctypes.byref(v), a, 0); print(name, v.value)
