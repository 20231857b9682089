OPENQASM 3.0;
include "stdgates.inc";
qubit[1] q;
gphase(0.25);
U(1.2, -0.4, 2.2) q[0];
