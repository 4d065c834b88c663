// Hand-written stand-in for the cmake-generated logger config: logging off, benchmarking off.
#ifndef KMI_REFCFG_BLISS_LOGGER_CONFIG_HPP
#define KMI_REFCFG_BLISS_LOGGER_CONFIG_HPP
#define USE_LOGGER 1
#define LOG_VERBOSITY 0
#endif
