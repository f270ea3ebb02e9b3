// Stand-in for <dlib/gui_widgets.h> on a headless build.  annonet_infer_main.cpp:22 includes the header but uses nothing from
// it; dlib's own header refuses to compile without X11 (DLIB_NO_GUI_SUPPORT -> #error).  shim/Makefile puts this directory in
// front of the reference's include path so that the reference's source file itself is compiled UNMODIFIED.
#pragma once
