"""ORACLE — test infrastructure only (see oracle/README.md).  Never imported by synthesizer_b200/."""
