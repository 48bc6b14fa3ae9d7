class _D:
    def __getattr__(self, n): return _D()
    def __call__(self, *a, **k): return _D()
import sys
class _M(type(sys)):
    def __getattr__(self, n): return _D()
sys.modules[__name__].__class__ = _M
