OPENQASM 3.0;
include "stdgates.inc";
qubit[2] q;
qubit[1] a;
ctrl @ h a[0], q[1];
ctrl @ ctrl @ z q[0], q[1], a[0];
y q[0];
