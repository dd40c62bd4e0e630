"""ORACLE package: CPU restatement of the reference's scoring arithmetic (test infrastructure only)."""
