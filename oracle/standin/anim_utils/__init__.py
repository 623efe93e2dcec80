"""Stand-in for eherr/anim_utils (named in reference README.md:6, not shipped,
not pinned).  TEST INFRASTRUCTURE ONLY; PARITY UNPINNED."""
