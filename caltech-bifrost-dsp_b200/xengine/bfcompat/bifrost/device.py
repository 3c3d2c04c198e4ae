"""``bifrost.device``: set_device / get_device / stream_synchronize."""
from xengine import runtime as _rt

set_device = _rt.set_device
get_device = _rt.get_device
stream_synchronize = _rt.stream_synchronize
