"""Stub of psana's ``Detector`` package (see ``tests/fake_psana/psana``): test infrastructure."""
