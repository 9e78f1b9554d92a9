// SizeMismatchException of libtimestep (reference: deps/libtimestep/include/libtimestep/exception/exception.h:11-24),
// thrown by the system constructor when x0, v0, theta0, omega0 differ in length
// (reference: libtimestep/rotational_system/rotational_system.h:69-70).
#ifndef LIBTIMESTEP_EXCEPTION_H
#define LIBTIMESTEP_EXCEPTION_H

#include <exception>
#include <string>
#include <utility>

class SizeMismatchException : public std::exception {
public:
    SizeMismatchException(std::string message, std::string file, int line)
        : text(std::move(message) + " (" + std::move(file) + ":" + std::to_string(line) + ")") {}
    const char* what() const noexcept override { return text.c_str(); }

private:
    std::string text;
};

// Any other failure of the CUDA backend (no device, out of memory, capacity): carries lg_last_error().
class Lgb200BackendError : public std::exception {
public:
    Lgb200BackendError(int code, std::string message) : code_(code), text(std::move(message)) {}
    const char* what() const noexcept override { return text.c_str(); }
    int code() const noexcept { return code_; }

private:
    int code_;
    std::string text;
};

#endif  // LIBTIMESTEP_EXCEPTION_H
