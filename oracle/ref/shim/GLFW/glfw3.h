// shim for GLFW (enginee/lib/init.sh:36): io.h / ui.h / render.h only need the opaque window type.  TEST INFRASTRUCTURE ONLY.
#pragma once
struct GLFWwindow;
