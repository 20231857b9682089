OPENQASM 3.0;
include "stdgates.inc";
qubit[2] q;
// a controlled global phase is a phase gate on its control: this is T on q[1] ...
ctrl @ gphase(0.7853981633974483) q[1];
// ... and a phase of pi/2 where q[0] = 0, i.e. on |00> and |01>
negctrl @ gphase(1.5707963267948966) q[0];
