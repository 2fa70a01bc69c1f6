"""oracle/ -- CPU restatements of the reference's algorithms.  TEST INFRASTRUCTURE ONLY:
imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs,
never by matchinglib_poselib_b200/."""
