"""Proposal NMS / top-k (reference lib/core/proposal.py) on the CUDA path."""
