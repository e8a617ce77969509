"""CPU oracle for the ORB front-end -- TEST INFRASTRUCTURE, not a product path.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this package (see DESIGN.md "Oracle").
"""
