/* oracle/ref_main.cpp -- TEST INFRASTRUCTURE ONLY.
 * Thin driver that calls the UNMODIFIED reference tool entry points
 * (normalize.h:29, decompose.h:29, decompose_blocksub.h, uniq.h:29, annotate_indels.h:29, view.h, index.h) compiled from /root/reference
 * by oracle/Makefile.  It replaces the reference's main.cpp (which drags in all ~60 tools) and
 * keeps its observable behaviour for these sub-commands: argv[1] selects the tool, and a
 * "Time elapsed: ..s" line is printed to stderr afterwards (main.cpp:90-111, 428-433). */
#include <cstdio>
#include <cstring>
#include <ctime>
#include <string>
#include "normalize.h"
#include "decompose.h"
#include "uniq.h"
#include "decompose_blocksub.h"
#include "annotate_indels.h"
#include "view.h"
#include "index.h"

int main(int argc, char** argv)
{
    clock_t t0 = clock();
    bool print = true;
    if (argc < 2)
    {
        fprintf(stderr, "usage: vt_ref normalize|decompose|decompose_blocksub|uniq|annotate_indels|view|index ...\n");
        return 1;
    }
    std::string cmd(argv[1]);
    if (cmd == "normalize") print = normalize(argc - 1, ++argv);
    else if (cmd == "decompose") decompose(argc - 1, ++argv);
    else if (cmd == "uniq") uniq(argc - 1, ++argv);
    else if (cmd == "decompose_blocksub") decompose_blocksub(argc - 1, ++argv);
    else if (cmd == "annotate_indels") annotate_indels(argc - 1, ++argv);
    else if (cmd == "view") print = view(argc - 1, ++argv);
    else if (cmd == "index") print = index(argc - 1, ++argv);
    else
    {
        fprintf(stderr, "Command not found: %s\n", argv[1]);
        return 1;
    }
    if (print)
    {
        double t = (double)(clock() - t0) / CLOCKS_PER_SEC;
        fprintf(stderr, "Time elapsed: %.2fs\n\n", t);
    }
    return 0;
}
