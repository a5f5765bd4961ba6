* hand-written: every RANGES case (E with +/- range, L, G), UP/LO/FX/MI/FR/BV bounds, negative RHS
NAME          RANGEBND
ROWS
 N  COST
 E  BAL1
 E  BAL2
 L  CAP1
 G  DEM1
 L  CAP2
 G  DEM2
COLUMNS
    XA        COST         2.5   BAL1         1
    XA        CAP1         1     DEM1         1
    XB        COST        -1     BAL1         1
    XB        BAL2         1     CAP2         2
    XC        COST         3     BAL2        -1
    XC        CAP1         1     DEM2         1
    XD        COST         0.5   BAL2         1
    XD        DEM1         1     CAP2        -1
    XE        COST        -2     CAP1         1
    XE        DEM2        -1     BAL1        -1
    XF        COST         1     DEM1         1
    XF        CAP2         1
RHS
    RHS       BAL1         4     BAL2        -3
    RHS       CAP1        10     DEM1         2
    RHS       CAP2         7     DEM2        -5
RANGES
    RNG       BAL1         2     BAL2        -1.5
    RNG       CAP1         6     DEM1         3.25
BOUNDS
 UP BND       XA           8
 LO BND       XB          -2
 UP BND       XB           5
 FX BND       XC           1.5
 MI BND       XD
 UP BND       XD           3
 FR BND       XE
 BV BND       XF
ENDATA
