// break_contigs_host.cpp -- main() of the drop-in break_contigs_start / break_contigs programs
// (-DPROGRAM_JOINS builds break_contigs).  Everything happens in salsa_break_contigs_run
// (csrc/cov_programs.cu), which replaces main() of break_contigs_start.cpp:271-463 and
// break_contigs.cpp:162-358.
#include "../../../include/salsa_b200.h"

int main(int argc, char **argv)
{
#ifdef PROGRAM_JOINS
    const int joins = 1;
#else
    const int joins = 0;
#endif
    return salsa_break_contigs_run(joins, argc - 1, argv + 1, 0, 0, 0);
}
