* hand-written: duplicated and linearly dependent equality rows after an L row (rank-deficient
* phase one with the artificial's index differing from its row), degenerate vertex
NAME DEGRED
ROWS
 N c
 L a_cap
 E e1
 E e2
 E e3
 G g1
 G g2
COLUMNS
 u c 1 a_cap 1
 u e1 1 e3 1
 u g1 1 g2 2
 v c 2 a_cap 1
 v e1 1 e2 1
 v e3 2 g1 1
 v g2 2
 w c -1 e2 1
 w e3 1 a_cap 1
 t c 0.25 e1 -1
 t e2 2 e3 1
 t g1 1 g2 2
RHS
 b a_cap 20 e1 4
 b e2 6 e3 10
 b g1 1 g2 2
ENDATA
