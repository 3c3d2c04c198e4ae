"""``bifrost.affinity``: set_core / get_core."""
from xengine import runtime as _rt

set_core = _rt.set_core
get_core = _rt.get_core
