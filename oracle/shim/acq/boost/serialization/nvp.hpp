// Test-infrastructure stub (NOT product code): Gnss_Synchro::serialize is a template that nothing here instantiates
#ifndef ORACLE_SHIM_ACQ_BOOST_NVP_H
#define ORACLE_SHIM_ACQ_BOOST_NVP_H
#define BOOST_SERIALIZATION_NVP(x) x
#endif
