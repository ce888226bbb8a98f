"""util -- host-side mirror of the reference's code/util.py (PrintingObject), the base class of every net.

Pure bookkeeping (a silence flag and a conditional print); nothing here touches the device.
"""


class PrintingObject:
    """Same surface as code/util.py:1-39: is_silent/get_silence/set_silence/unset_silence/with_silence/
    silence()/_print()."""

    class SilenceSignal:
        """Context manager returned by silence(): temporarily overrides the flag (code/util.py:3-12)."""

        def __init__(self, obj, value):
            self.obj, self.new_silent, self.old_silent = obj, value, None

        def __enter__(self):
            self.old_silent = self.obj.get_silence()
            self.obj.set_silence(self.new_silent)

        def __exit__(self, exception_type, exception_value, traceback):
            self.obj.set_silence(self.old_silent)

    def __init__(self):
        self.silent = True

    def is_silent(self):
        return self.silent

    get_silence = is_silent

    def set_silence(self, value=True):
        self.silent = value
        return self

    def unset_silence(self):
        return self.set_silence(False)

    with_silence = set_silence

    def silence(self, value=True):
        return PrintingObject.SilenceSignal(self, value)

    def _print(self, *args, **kwargs):
        if not self.silent:
            print(*args, **kwargs)
