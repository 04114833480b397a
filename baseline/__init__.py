"""Reference arm of bench.py: the reference's own CPU code for the path, installed (copied unmodified) under
``baseline/_ref`` by ``__graft_entry__.build()``.  Nothing here is part of the product."""
