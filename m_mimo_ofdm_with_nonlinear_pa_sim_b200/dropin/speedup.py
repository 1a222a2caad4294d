"""Same name as the reference's speedup.py (an optional-numba shim): the numeric work of this
package runs in CUDA kernels, so `jit` is the identity decorator here."""


def jit(pyfunc=None, **kwargs):
    def wrap(func):
        return func
    if pyfunc is not None:
        return wrap(pyfunc)
    return wrap
