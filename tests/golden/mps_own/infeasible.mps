* hand-written: infeasible
NAME INFEAS
ROWS
 N c
 L r1
 G r2
COLUMNS
 x c 1 r1 1
 x r2 1
 y c 1 r1 1
 y r2 1
RHS
 b r1 3 r2 5
ENDATA
