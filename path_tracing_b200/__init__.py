"""B200-native bidirectional path tracing (the cpu_bdpt.cpp sample loop of HongMJ1315/Path_Tracing) behind a C ABI."""
from . import _abi  # noqa: F401
