from ddsp_pytorch_b200.modules import Reverb  # noqa: F401  (replaces model/ddsp/reverb.py)
