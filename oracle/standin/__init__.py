"""Stand-in for `anim_utils` and `transformations` (absent from /root/reference).

Put this directory on sys.path (oracle.refload does) so that the reference's
unmodified modules can `import anim_utils...` / `import transformations`.
PARITY UNPINNED: semantics follow SURVEY.md Appendix A.
"""
