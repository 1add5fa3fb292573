"""CPU oracle of the PagedAttention decode path -- test infrastructure only.

Import from tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs;
never from the product package.
"""
