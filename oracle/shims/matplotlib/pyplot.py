"""TEST INFRASTRUCTURE ONLY -- headless stand-in: subplots() returns inert objects so that the reference's
analysis/report.py functions can run to their return statements (their numbers are what is recorded)."""


class _Inert(object):
    def __getattr__(self, name):
        return _Inert()

    def __call__(self, *a, **k):
        return _Inert()

    def __iter__(self):
        return iter(())


def subplots(*a, **k):
    return _Inert(), _Inert()


def show(*a, **k):
    return None
