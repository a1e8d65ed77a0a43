// The GNU Radio faces of the two B200 blocks: thin gr::block subclasses whose general_work forwards to the block objects of this
// tree, with the reference's scheduler contract and message ports
//   dll_pll_veml_tracking.cc:85-152 (io signatures, set_relative_rate, set_max_noutput_items(1), "events" out, "telemetry_to_trk" in),
//   :607-614 (forecast: 2 * vector_length), :617-641 (telemetry fault -> forced loss of lock), :1023 (event 3 = loss of lock);
//   pcps_acquisition.cc:50-52 (io signatures), :362,389 (events 1 / 2), :902-1031 (general_work: consume what the FSM took).
// Inside GNSS-SDR these are what get_left_block() / get_right_block() of the adapters return (build with -DGB200_WITH_GNURADIO: the real
// <gnuradio/block.h> and <pmt/pmt.h>); without GNU Radio the same code compiles against gnuradio_stub/ (declarations of the members used
// here, plus a one-block driver for the tests).
#ifndef GB200_GR_BLOCKS_H
#define GB200_GR_BLOCKS_H

#include "b200_dll_pll_veml_tracking.h"
#include "b200_e5a_noncoherent_iq_acquisition_caf.h"
#include "b200_pcps_acquisition.h"
#include "b200_pcps_acquisition_fine_doppler.h"
#include "b200_pcps_quicksync_acquisition.h"
#include <gnuradio/block.h>
#include <memory>

namespace gb200h
{
class b200_tracking_gr : public gr::block
{
public:
    b200_tracking_gr(std::shared_ptr<b200_dll_pll_veml_tracking> impl, size_t item_size, unsigned vector_length);
    void forecast(int noutput_items, gr_vector_int& ninput_items_required) override;
    int general_work(int noutput_items, gr_vector_int& ninput_items, gr_vector_const_void_star& input_items, gr_vector_void_star& output_items) override;
    std::shared_ptr<b200_dll_pll_veml_tracking> impl() { return d_impl; }

private:
    void msg_handler_telemetry_to_trk(const pmt::pmt_t& msg);
    std::shared_ptr<b200_dll_pll_veml_tracking> d_impl;
    unsigned d_vector_length;
};

class b200_acquisition_gr : public gr::block
{
public:
    b200_acquisition_gr(std::shared_ptr<b200_pcps_acquisition> impl, size_t item_size);
    int general_work(int noutput_items, gr_vector_int& ninput_items, gr_vector_const_void_star& input_items, gr_vector_void_star& output_items) override;
    std::shared_ptr<b200_pcps_acquisition> impl() { return d_impl; }

private:
    std::shared_ptr<b200_pcps_acquisition> d_impl;
};

// The other acquisition strategies, same pattern: "events" out port (1 positive / 2 negative), consume_each of what the FSM took.
//   pcps_acquisition_fine_doppler_cc.cc:40-68 (gr_complex in, Gnss_Synchro out), :171-178 (forecast: fft_size items), :450-563
class b200_fine_doppler_acquisition_gr : public gr::block
{
public:
    explicit b200_fine_doppler_acquisition_gr(std::shared_ptr<b200_pcps_acquisition_fine_doppler> impl);
    void forecast(int noutput_items, gr_vector_int& ninput_items_required) override;
    int general_work(int noutput_items, gr_vector_int& ninput_items, gr_vector_const_void_star& input_items, gr_vector_void_star& output_items) override;

private:
    std::shared_ptr<b200_pcps_acquisition_fine_doppler> d_impl;
};

//   galileo_e5a_noncoherent_iq_acquisition_caf_cc.cc:55-75, :295-777
class b200_e5a_caf_acquisition_gr : public gr::block
{
public:
    explicit b200_e5a_caf_acquisition_gr(std::shared_ptr<b200_e5a_noncoherent_iq_acquisition_caf> impl);
    int general_work(int noutput_items, gr_vector_int& ninput_items, gr_vector_const_void_star& input_items, gr_vector_void_star& output_items) override;

private:
    std::shared_ptr<b200_e5a_noncoherent_iq_acquisition_caf> d_impl;
};

//   pcps_quicksync_acquisition_cc.cc:54-67 (VECTOR items in), :214-550.  The reference's item is sampled_ms * samples_per_ms gr_complex; here the item is
//   what its adapters' stream_to_vector makes (b200_pcps_quicksync_acquisition::item_length)
class b200_quicksync_acquisition_gr : public gr::block
{
public:
    explicit b200_quicksync_acquisition_gr(std::shared_ptr<b200_pcps_quicksync_acquisition> impl);
    int general_work(int noutput_items, gr_vector_int& ninput_items, gr_vector_const_void_star& input_items, gr_vector_void_star& output_items) override;

private:
    std::shared_ptr<b200_pcps_quicksync_acquisition> d_impl;
};

std::shared_ptr<b200_tracking_gr> make_b200_tracking_gr(std::shared_ptr<b200_dll_pll_veml_tracking> impl, size_t item_size, unsigned vector_length);
std::shared_ptr<b200_acquisition_gr> make_b200_acquisition_gr(std::shared_ptr<b200_pcps_acquisition> impl, size_t item_size);
}  // namespace gb200h
#endif
