OPENQASM 3.0;
include "stdgates.inc";
qubit[3] q;
// X on q[2] when q[0] = 1 and q[1] = 0
ctrl @ negctrl @ x q[0], q[1], q[2];
