"""iu-b200: batched point location and interpolation on unstructured grids, native to NVIDIA B200.

The compute path is hand-written CUDA for sm_100a inside ``csrc/`` behind a C ABI
(``include/iu_b200.h`` -> ``libiu_b200.so``).  ``api`` mirrors the reference's ``iu_*`` Fortran
interface (m_interp_unstructured.f90:77-101) on top of that ABI via ctypes.  There is no CPU
fallback: importing ``api`` and creating a grid fails loudly when the CUDA library is missing.
"""
__version__ = "0.1.0"
