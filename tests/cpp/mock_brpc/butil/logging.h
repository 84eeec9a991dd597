// TEST DOUBLE of butil/logging.h: LOG(severity) << ... goes to stderr.
#pragma once
#include <iostream>
#define LOG(severity) std::cerr << "[" #severity "] "
