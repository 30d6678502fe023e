"""hddm_b200 -- B200-native (sm_100a) WFPT likelihood and sequential-sampling simulators behind HDDM's own API.

Importing this package is cheap and touches neither CUDA nor the shared library (models are pickled/unpickled by
the reference, /root/reference/hddm/models/base.py:1394-1403, which re-imports the likelihood module).  The library
is loaded on first use; there is no CPU fallback: a missing library or GPU raises.
"""
__version__ = "0.1.0"
