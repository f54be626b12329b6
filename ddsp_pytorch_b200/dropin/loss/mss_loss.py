from ddsp_pytorch_b200.modules import MSSLoss, SSSLoss  # noqa: F401  (replaces loss/mss_loss.py)
