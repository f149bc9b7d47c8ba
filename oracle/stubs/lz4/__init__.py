"""Stub of the lz4 package (test infrastructure only): identity codec so the
reference's file_helper imports. No arithmetic on the hot path touches it."""
