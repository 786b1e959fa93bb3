/* vtn_tools.h -- the reference's TOOL entry points for the decompose / normalize path, exported from libvtn_tools.so (C++ host
 * library over the C ABI of vtn.h; vt_b200/csrc/vtn_tools.cpp).  A tool takes the argument vector the reference's tool takes
 * (argv[0] = the sub-command name, then vt's own flags), prints the same option / statistics text on stderr and writes the same VCF;
 * the per-record arithmetic runs in libvtn's CUDA kernels.  Like the reference's tools they end the process with exit(1) on a fatal
 * input error (normalize.cpp:209-210, variant_manip.cpp:72-74); they have no CPU fallback: without a CUDA device they exit 1.
 *
 *   this library                      replaces in the reference                                  called from
 *   vtn_tool_normalize                bool normalize(int argc, char** argv)        normalize.h:29   main.cpp:393-396
 *   vtn_tool_decompose                void decompose(int argc, char** argv)        decompose.h:29   main.cpp:193-196
 *   vtn_tool_annotate_indels          void annotate_indels(int argc, char** argv)  annotate_indels.h:29  main.cpp (annotate_indels branch)
 *   vtn_tool_uniq                     void uniq(int argc, char** argv)             uniq.h:29
 *   vtn_tool_decompose_blocksub       void decompose_blocksub(int argc, char**)    decompose_blocksub.h:29
 *   vtn_tool_main                     main()'s dispatch + "Time elapsed" line      main.cpp:145-146, 428-433
 */
#ifndef VTN_TOOLS_H
#define VTN_TOOLS_H

#ifdef __cplusplus
extern "C" {
#endif

/* returns non-zero when the caller should print the elapsed time (the reference returns `print`, normalize.cpp:391-400) */
int vtn_tool_normalize(int argc, char** argv);
void vtn_tool_decompose(int argc, char** argv);
void vtn_tool_annotate_indels(int argc, char** argv);
void vtn_tool_uniq(int argc, char** argv);
void vtn_tool_decompose_blocksub(int argc, char** argv);

/* argv as the program got it: argv[1] = sub-command.  Returns the process exit status. */
int vtn_tool_main(int argc, char** argv);

/* on != 0: once a tool's output file is complete and closed, vtn_tool_main ends the process (_exit) instead of releasing the CUDA
 * context and the page-locked buffers one by one -- what the command-line program `vtn` does.  Off by default, so a host program
 * that calls a tool gets every resource back when the tool returns. */
void vtn_tools_end_process_after_output(int on);

#ifdef __cplusplus
}
#endif
#endif
