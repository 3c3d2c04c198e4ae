"""``bifrost.proclog.ProcLog``."""
from xengine.runtime import ProcLog  # noqa: F401
