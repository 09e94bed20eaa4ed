// run_blt.cpp -- the stand-alone solver, same command line and output as the reference (libblt/run_blt.cc:70-198):
//   run_blt [-m] [-v] [-h] [-y] <file.bilp>
//   PROBLEM <file> RUNTIME <s> niter XXX npru XXX rc XXX check  SAT | UNSAT |INPUT_ERROR
//   PROBLEM <file> MODEL (x0, x1, ...)                 (with -m, when SAT)
// exit status = the blt_check() code.  Set BLT_STATS=1 to get the real counters on stderr (the drop-in keeps XXX).
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <time.h>
#include <unistd.h>

#include <fstream>
#include <iostream>
#include <string>
#include <vector>

#include "../../../include/blt.h"
#include "../../../include/codes.h"
#include "context.h"


static void print_usage(char *argv[]) {
    std::cout << "Usage: " << argv[0] << " [options]  <input>\n\n  [options]\n"
              << "  -v          print version information\n  -h          print this help message\n"
              << "  -m          print out a model if SAT\n"
              << "  -y          enable \"sound and complete mode\" using the Yices SMT solver\n"
              << "              (libyices shared library must be in the runtime linker's search path)\n\n"
              << "  <input>     input filename (BILP format)\n" << std::endl;
}

int main(int argc, char *argv[]) {
    // Extension: run_blt [-m] [-y] --latcon <file.lat> <file.con> solves a legacy lattice/constraint pair the way the
    // reference's regression driver does (solve_raw, libblt/test_regression.cc:36-61); same PROBLEM line, same exit status.
    for (int i = 1; i < argc; ++i) {
        if (std::string(argv[i]) != "--latcon") continue;
        if (i + 2 >= argc) { std::cerr << "Error: --latcon needs <file.lat> <file.con>" << std::endl; return 1; }
        bool model = false, sound = false;
        for (int j = 1; j < i; ++j) { if (std::string(argv[j]) == "-m") model = true; if (std::string(argv[j]) == "-y") sound = true; }
        std::vector<int64> x(4096);
        int rows = 0;
        uint64_t ni = 0, np = 0;
        const clock_t start = clock();
        const int rc = blt_b200_solve_raw_files(argv[i + 1], argv[i + 2], sound ? 1 : 0, x.data(), (int)x.size(), &rows, &ni, &np);
        const double elapsed = (1.0 * (clock() - start)) / CLOCKS_PER_SEC;
        std::cout << "PROBLEM " << argv[i + 1] << " RUNTIME " << elapsed << " "
                  << "niter " << "XXX" << " npru " << "XXX" << " " << "rc " << "XXX" << " check "
                  << (rc == BLT_CHECK_SAT ? " SAT " : (rc == BLT_CHECK_UNSAT ? " UNSAT " : "INPUT_ERROR")) << std::endl;
        if (const char *e = getenv("BLT_STATS")) if (e[0] == '1') std::cerr << "BLT_STATS niter " << ni << " nprune " << np << " rc " << rc << std::endl;
        if (model && rc == BLT_CHECK_SAT) {
            std::cout << "PROBLEM " << argv[i + 1] << " MODEL (";
            for (int k = 0; k < rows; ++k) std::cout << x[k] << (k < rows - 1 ? ", " : ")\n");
        }
        return rc;
    }
    bool do_model = false, do_version = false, do_yices = false;
    int opt;
    while ((opt = getopt(argc, argv, "mhvy")) != -1) {
        switch (opt) {
            case 'm': do_model = true; break;
            case 'v': do_version = true; break;
            case 'y': do_yices = true; break;
            default: print_usage(argv); return 1;
        }
    }
    if (do_version) { std::cout << "BLT version " << __BLT_VERSION << std::endl; return 0; }
    if (argc - optind != 1) {
        std::cerr << "Error: Invalid numer of non-option arguments!" << std::endl;
        print_usage(argv);
        return 1;
    }
    const std::string fn(argv[optind]);
    std::ifstream infile(fn);
    std::vector<int64> M, C;
    size_t m = 0, n = 0;  // m inequality rows, n variables
    if (read_bilp(infile, M, C, m, n) != 0) { std::cerr << "Error: cannot parse input file: " << fn << std::endl; return 1; }

    CContext ctx = blt_init(do_yices);
    if (!ctx) return BLT_STATUS_ERROR;
    const size_t var_name_size = 2 + lround(log10((double)n) + 0.5);
    std::vector<char> var_name(var_name_size + 1);
    std::vector<CExpr> var(n);
    for (size_t i = 0; i < n; ++i) {
        snprintf(var_name.data(), var_name_size, "%zu", i);
        var[i] = blt_var(ctx, var_name.data());
    }
    for (size_t i = 0; i < m; ++i) {
        CExpr e = blt_const(ctx, 0, 1);
        for (size_t j = 0; j < n; ++j) e = blt_add(ctx, e, blt_smul(ctx, M[i * n + j], 1, var[j]));
        (void)blt_assume(ctx, C[2 * i], 1, e, C[2 * i + 1], 1);
    }
    const clock_t start = clock();
    const int rc = blt_check(ctx);
    const double elapsed = (1.0 * (clock() - start)) / CLOCKS_PER_SEC;
    std::cout << "PROBLEM " << fn << " RUNTIME " << elapsed << " "
              << "niter " << "XXX" << " npru " << "XXX" << " " << "rc " << "XXX" << " check "
              << (rc == BLT_CHECK_SAT ? " SAT " : (rc == BLT_CHECK_UNSAT ? " UNSAT " : "INPUT_ERROR")) << std::endl;
    if (const char *e = getenv("BLT_STATS")) {
        if (e[0] == '1') {
            uint64_t ni = 0, np = 0; int rank = 0; double lm = 0, sm = 0;
            blt_b200_last_stats(ctx, &ni, &np, &rank, &lm, &sm);
            std::cerr << "BLT_STATS niter " << ni << " nprune " << np << " rank " << rank << " lll_ms " << lm
                      << " search_ms " << sm << " rc " << rc << std::endl;
        }
    }
    if (do_model && rc == BLT_CHECK_SAT) {
        std::cout << "PROBLEM " << fn << " MODEL (";
        for (size_t i = 0; i < n; ++i) {
            int64 x = 0;
            snprintf(var_name.data(), var_name_size, "%zu", i);
            blt_model(ctx, var_name.data(), &x);
            std::cout << x << (i < n - 1 ? ", " : ")\n");
        }
    }
    return rc;
}
