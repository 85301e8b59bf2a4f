// Oracle shim (test infrastructure only).  The reference includes "Uncertainty.h"
// (Constraints.h:17, member at Constraints.h:155, call at Constraints.cpp:188-201) but ships no
// such file, so the grid-map uncertainty cost term cannot be pinned.  This stub returns zeros and
// the harness always calls clear_uncertainty_map() so the term is never evaluated.
#pragma once
#include "struct_x_vx_mx_vmx.h"
class Uncertainty
{
public:
    Uncertainty() {}
    struct_x_vx_mx_vmx get_uncertainty_cost(const Eigen::VectorXd &)
    {
        struct_x_vx_mx_vmx r;
        r.x = 0.0;
        r.vx = Eigen::VectorXd::Zero(4);
        r.mx = Eigen::MatrixXd::Zero(4, 4);
        return r;
    }
};
