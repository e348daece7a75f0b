"""ORACLE — test infrastructure only.  See oracle/oracle_math.h for the header statement."""
