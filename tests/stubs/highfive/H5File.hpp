// empty stand-in: the driver syntax check (tests/test_md_shim.py) only compares diagnostics between header sets
