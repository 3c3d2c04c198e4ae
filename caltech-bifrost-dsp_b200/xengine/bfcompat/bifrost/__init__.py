"""Stand-in for the names the reference's X-engine blocks import from bifrost (see ../README.md).  NOT bifrost."""
from . import ndarray as _ndarray_module
from .ndarray import ndarray  # noqa: F401   `import bifrost.ndarray as BFArray` must yield the CLASS, as with bifrost
from . import affinity, device, libbifrost, linalg, proclog, ring, unpack  # noqa: F401
from .map import map  # noqa: F401            `from bifrost import map as BFMap`
