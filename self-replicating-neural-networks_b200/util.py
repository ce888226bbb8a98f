"""util -- host-side stand-in for the reference's code/util.py.

The reference's only utility is `PrintingObject`, the base class of every net: a verbosity switch plus a
print that honours it.  Nothing here touches the device; it exists so that user code written against the
reference (`net.silence()`, `net.with_silence(False)`, `net._print(...)`) keeps working.
"""
import contextlib


class PrintingObject:
    """Verbosity switch shared by all nets.  Objects start silent, exactly like the reference's
    (code/util.py:14-15); every setter returns `self` so calls can be chained."""

    def __init__(self):
        self.silent = True

    # -- queries -------------------------------------------------------------------------------------
    def is_silent(self):
        return bool(self.silent)

    def get_silence(self):
        return self.is_silent()

    # -- setters (chainable) -------------------------------------------------------------------------
    def set_silence(self, value=True):
        self.silent = value
        return self

    def unset_silence(self):
        return self.set_silence(False)

    def with_silence(self, value=True):
        return self.set_silence(value)

    # -- scoped override -----------------------------------------------------------------------------
    class SilenceSignal(contextlib.AbstractContextManager):
        """`with net.silence(False): ...` -- the flag is switched for the block and put back afterwards,
        also when the block raises."""

        def __init__(self, obj, value):
            self.obj, self.new_silent, self.old_silent = obj, value, obj.get_silence()

        def __enter__(self):
            self.old_silent = self.obj.get_silence()
            self.obj.set_silence(self.new_silent)
            return self

        def __exit__(self, exc_type, exc, tb):
            self.obj.set_silence(self.old_silent)
            return False

    def silence(self, value=True):
        return type(self).SilenceSignal(self, value)

    # -- output --------------------------------------------------------------------------------------
    def _print(self, *args, **kwargs):
        """print(...) unless the object is silent."""
        if self.is_silent():
            return
        print(*args, **kwargs)
