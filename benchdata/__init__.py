"""benchdata -- the synthetic NA12878-like workload of SURVEY.md 8(d) (bench and test support;
not part of the product library)."""
