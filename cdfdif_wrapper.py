"""Top-level `cdfdif_wrapper` module: the name the reference imports (`import cdfdif_wrapper as cdfdif`,
/root/reference/hddm/__init__.py:24,31-32).  Everything lives in hddm_b200.cdfdif."""
from hddm_b200.cdfdif import dmat_cdf_array  # noqa: F401
