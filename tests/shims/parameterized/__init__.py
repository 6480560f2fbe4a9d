# minimal stand-in for the `parameterized` package (not installed here): parameterized.expand only
import functools, inspect, sys
class parameterized:
    @staticmethod
    def expand(cases, **kw):
        def deco(fn):
            frame = sys._getframe(1).f_locals
            for i, case in enumerate(cases):
                args = case if isinstance(case, (tuple, list)) else (case,)
                def make(a):
                    @functools.wraps(fn)
                    def t(self): return fn(self, *a)
                    return t
                t = make(tuple(args)); t.__name__ = f"{fn.__name__}_{i}"
                frame[t.__name__] = t
            return None
        return deco
