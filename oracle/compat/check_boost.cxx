#include <boost/program_options.hpp>
int main(){return 0;}
