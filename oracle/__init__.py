"""CPU oracle for the Mercury importance-sampling hot path (test infrastructure only)."""
