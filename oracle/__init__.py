"""CPU oracle for the deep-go move-prediction hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``oracle/`` is part of the product: only
``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline / reference arm
may import it, and only as the checker or the timed CPU baseline.  The product path
(``deep-go_b200/`` + ``libdeepgo_b200.so``) never imports or links anything from here.

Parity status:
  * board analysis (makedata.lua) ........ PINNED by the reference's 4,398 fixture records
  * record format (Torch7 binary) ........ PINNED by parsing the same files
  * preprocess / net / NLL / SGD ......... parity UNPINNED (Torch7 cannot run here and the
    reference has no tests for them); the restatement follows the cited lines.
  * symmetry ids 1..7 .................... parity UNPINNED (reference transform() is identity)
"""
