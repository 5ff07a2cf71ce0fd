"""ORACLE — test infrastructure only.  See seva_oracle.py."""
