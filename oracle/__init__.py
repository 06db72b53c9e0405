"""TEST INFRASTRUCTURE ONLY -- CPU oracle for the generation path.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import anything from this package. The product package never does (tests/test_no_oracle_in_product.py
enforces it)."""
