"""
custom_porechop_abi_b200 -- B200-native drop-in for Porechop_ABI's adapter-alignment hot path.

Only what that path needs lives here: csrc/ (sm_100a kernels + the C ABI, built into cpp_functions.so),
cpp_function_wrappers.py (the reference's ctypes boundary, same names) and engine.py (batched submit/collect).
Importing the package does not load the CUDA library; importing cpp_function_wrappers / engine does, and
fails hard when it is missing, as the reference does (porechop_abi/cpp_function_wrappers.py:21-25).
"""
__version__ = '0.1.0'
