"""pytest-side wrapper of baseline/ref_loader.py: skip (with the reason) when baseline/_ref was not shipped."""
import pytest

from baseline.ref_loader import REF, have_ref, load as _load  # noqa: F401


def load():
    if not have_ref():
        pytest.skip("baseline/_ref is missing: run `python baseline/build_ref.py` in the build container")
    return _load()
