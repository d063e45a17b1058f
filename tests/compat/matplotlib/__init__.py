"""Inert stand-in for matplotlib (orchestrator.py:19, 398-415): the confusion-matrix / ROC PNGs are outside the hot path."""
__version__ = "0.0-shim"
