// phantom-fhe "example.h" stand-in (dotprod_test.cu:13): print_parameters / print_vector.  The text is the interface: it is
// what reliability_test/data/bits1-16_num1.txt:4-12 records for print_parameters and :17-18 for print_vector(v, 3, 7).
#pragma once
#include <algorithm>
#include <iomanip>
#include <iostream>
#include <vector>
#include "phantom.h"

inline void print_parameters(const PhantomContext& context) {
    const auto& parms = context.parms();
    const char* name = parms.scheme() == phantom::scheme_type::bgv ? "BGV" : parms.scheme() == phantom::scheme_type::bfv ? "BFV" : "CKKS";
    std::cout << "/" << std::endl;
    std::cout << "| Encryption parameters :" << std::endl;
    std::cout << "|   scheme: " << name << std::endl;
    std::cout << "|   poly_modulus_degree: " << parms.poly_modulus_degree() << std::endl;
    int total = 0;
    for (const auto& m : parms.coeff_modulus()) total += m.bit_count();
    std::cout << "|   coeff_modulus size: " << total << " (";
    for (size_t i = 0; i < parms.coeff_modulus().size(); i++) std::cout << (i ? " + " : "") << parms.coeff_modulus()[i].bit_count();
    std::cout << ") bits" << std::endl << std::endl;
    for (const auto& m : parms.coeff_modulus()) std::cout << m.value() << " ,  ";
    std::cout << std::endl << std::endl;
    std::cout << "\\" << std::endl;
}

template <typename T>
inline void print_vector(std::vector<T> vec, std::size_t print_size = 4, int prec = 3) {
    std::ios old_fmt(nullptr);
    old_fmt.copyfmt(std::cout);
    const std::size_t slot_count = vec.size();
    std::cout << std::fixed << std::setprecision(prec) << std::endl;
    if (slot_count <= 2 * print_size) {
        std::cout << "    [";
        for (std::size_t i = 0; i < slot_count; i++) std::cout << " " << vec[i] << ((i != slot_count - 1) ? "," : " ]\n");
    } else {
        std::cout << "    [";
        for (std::size_t i = 0; i < print_size; i++) std::cout << " " << vec[i] << ",";
        std::cout << " ...,";
        for (std::size_t i = slot_count - print_size; i < slot_count; i++) std::cout << " " << vec[i] << ((i != slot_count - 1) ? "," : " ]\n");
    }
    std::cout << std::endl;
    std::cout.copyfmt(old_fmt);
}
