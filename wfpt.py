"""Top-level `wfpt` module: the name the reference imports (`import wfpt`, /root/reference/hddm/__init__.py:29;
`from wfpt import wiener_like_rlddm`, hddm/models/hddm_rl.py:12).  Everything lives in hddm_b200.wfpt."""
from hddm_b200.wfpt import *  # noqa: F401,F403
from hddm_b200.wfpt import FAITHFUL, FAST, WfptError, copy, np  # noqa: F401
