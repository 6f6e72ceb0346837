"""``lib`` package shim: lets the reference's scripts (train.py:15, eval_pf_pascal.py:10-15, eval_inloc.py:16-19) run
unchanged on the B200 head.

Put this directory FIRST on PYTHONPATH and the reference checkout after it:

    PYTHONPATH=/path/to/repo/ncnet_b200/dropin:/path/to/repo:/path/to/ncnet python train.py ...

``lib.model``, ``lib.conv4d``, ``lib.point_tnf`` and ``lib.eval_util`` then resolve to the modules here (re-exports of
``ncnet_b200``); every other ``lib.*`` module (datasets, transforms, torch_util, ...) resolves to the reference's
own file, because the reference's ``lib`` directory is appended to this package's search path.  Set
``NCNET_REFERENCE_ROOT`` if the reference checkout is not on ``sys.path``.
"""
import os
import sys

_here = os.path.dirname(os.path.abspath(__file__))
_roots = [os.environ.get("NCNET_REFERENCE_ROOT")] + list(sys.path)
for _root in _roots:
    if not _root:
        continue
    _cand = os.path.join(os.path.abspath(_root), "lib")
    if _cand != _here and os.path.isfile(os.path.join(_cand, "torch_util.py")) and _cand not in __path__:
        __path__.append(_cand)
        break
