"""Where does process start-up go?  cuInit / context, first engine, later engines."""
import ctypes, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
t0 = time.perf_counter()
from pt_b200.engine import Engine, load_library
L = load_library()
t1 = time.perf_counter()
cuda = ctypes.CDLL("libcudart.so.12")
cuda.cudaFree(0)
t2 = time.perf_counter()
es = []
ts = []
for i in range(4):
    a = time.perf_counter()
    es.append(Engine(27, 100, 1949, 4, 51, 100))
    ts.append(time.perf_counter() - a)
print(f"dlopen {t1-t0:.3f}s  cudaFree(0) (context) {t2-t1:.3f}s  engines {[round(x,4) for x in ts]}")
