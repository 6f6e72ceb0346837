"""lib.eval_util -> ncnet_b200.eval_util (pck, pck_metric of the reference's lib/eval_util.py)."""
from ncnet_b200.eval_util import pck, pck_metric  # noqa: F401
