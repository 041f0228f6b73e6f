"""ORACLE package: CPU restatement of the reference's caption-decoder path.
Test infrastructure only -- see oracle/restated.py header for who may import it."""
