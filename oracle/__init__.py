"""TEST INFRASTRUCTURE.  CPU oracle of the per-pixel hot path; see oracle/rt_oracle.c.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / reference arm may import
this package.  The product (real_time_ray_tracer_b200) never does."""
