OPENQASM 3.0;
include "stdgates.inc";
qubit[2] q;
rx(0.3) q[0];
ry(-1.1) q[1];
rz(2.0) q[0];
negctrl @ rz(0.5) q[1], q[0];
sdg q[1];
tdg q[0];
