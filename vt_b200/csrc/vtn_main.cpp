// vtn -- the command-line program: vt's main.cpp dispatch (main.cpp:193-196, 221-224, 393-396, 428-433) over libvtn_tools.so
#include "../../include/vtn_tools.h"

int main(int argc, char** argv)
{
    vtn_tools_end_process_after_output(1);
    return vtn_tool_main(argc, argv);
}
