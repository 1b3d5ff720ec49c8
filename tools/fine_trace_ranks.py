"""Critical-path marks (BX_FINE_TRACE build) of the partitioned fine-mesh kernel, R ranks on one device."""
import ctypes, os, sys, runpy
lib = os.path.abspath(os.environ["BX_VARIANT_LIB"])
sys.argv = ["finemesh_ranks_local.py"] + sys.argv[1:]
runpy.run_path(os.path.join(os.path.dirname(__file__), "finemesh_ranks_local.py"), run_name="__main__")
raw = ctypes.CDLL(lib)
buf = (ctypes.c_ulonglong * 32)()
raw.bx_fine_trace_read(buf, 1)
cyc = sum(buf[i] for i in range(16))
for i in range(16):
    if buf[16 + i]:
        print(f"  [{i:2d}] {100.0 * buf[i] / cyc:5.1f} %  {buf[i] / buf[16 + i]:8.0f} cycles x {buf[16 + i]}")
