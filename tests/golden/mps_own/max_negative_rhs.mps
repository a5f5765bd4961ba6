* hand-written: OBJSENSE MAX, negative right-hand sides on L and G rows, upper-only variable
NAME MAXNEG
OBJSENSE
    MAX
ROWS
 N obj
 L r1
 G r2
 L r3
 E r4
COLUMNS
 x obj 3 r1 1
 x r2 -1 r4 1
 y obj 2 r1 1
 y r3 -1 r4 -1
 z obj -1 r2 1
 z r3 1 r4 2
RHS
 rhs r1 12 r2 -8
 rhs r3 -1 r4 3
BOUNDS
 UP bnd z 4
 MI bnd z
 UP bnd y 9
ENDATA
