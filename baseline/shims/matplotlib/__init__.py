"""Import shim: matplotlib is imported by the reference preprocessors.py, used only under __main__."""
