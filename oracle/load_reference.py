"""Import the UNMODIFIED reference hot-path classes from /root/reference (build container only).

TEST INFRASTRUCTURE ONLY (see oracle/geotx_oracle.py header).  /root/reference does not
exist on the GPU box; callers must check ``reference_available()`` first.

Plain ``import geotransformer.modules.geotransformer`` fails in this image for reasons that
have nothing to do with the hot path: the CPU C++ ext is not built, and
open3d / IPython / ipdb are not installed.  None of them is used by the path, so they are
replaced by empty stub modules before the import.
"""
import os
import sys
import types

REFERENCE_ROOT = os.environ.get('GEOTX_REFERENCE_ROOT', '/root/reference')


def reference_available():
    return os.path.isdir(os.path.join(REFERENCE_ROOT, 'geotransformer', 'modules'))


def load_reference():
    """Returns the reference's (GeometricTransformer, GeometricStructureEmbedding) classes."""
    if not reference_available():
        raise RuntimeError(f'reference tree not found at {REFERENCE_ROOT}')
    sys.dont_write_bytecode = True  # the reference tree is read-only
    for name in ('IPython', 'geotransformer.ext', 'open3d', 'ipdb'):
        if name not in sys.modules:
            mod = types.ModuleType(name)
            if name == 'IPython':
                mod.embed = lambda *a, **k: None
            sys.modules[name] = mod
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    from geotransformer.modules.geotransformer.geotransformer import (  # noqa: E402
        GeometricStructureEmbedding, GeometricTransformer)
    return GeometricTransformer, GeometricStructureEmbedding
