#include "b200_gr_blocks.h"
#ifdef GB200_WITH_GNURADIO
#include <gnuradio/io_signature.h>
#include <boost/any.hpp>
#endif

namespace gb200h
{
// ------------------------------------------------------------------ tracking (dll_pll_veml_tracking.cc:85-152)
b200_tracking_gr::b200_tracking_gr(std::shared_ptr<b200_dll_pll_veml_tracking> impl, size_t item_size, unsigned vector_length)
    : gr::block("b200_dll_pll_veml_tracking", gr::io_signature::make(1, 1, static_cast<int>(item_size)), gr::io_signature::make(1, 1, sizeof(Gnss_Synchro))),
      d_impl(std::move(impl)),
      d_vector_length(vector_length)
{
    this->set_relative_rate(1, static_cast<uint64_t>(d_vector_length));  // one Gnss_Synchro per code period at most (:129-133)
    this->set_max_noutput_items(1);                                      // prevent telemetry symbols accumulating in output buffers (:135)
    this->message_port_register_out(pmt::mp("events"));                  // :138
    this->message_port_register_in(pmt::mp("telemetry_to_trk"));         // :141
    this->set_msg_handler(pmt::mp("telemetry_to_trk"), [this](const pmt::pmt_t& msg) { msg_handler_telemetry_to_trk(msg); });
    d_impl->set_event_sink([this](int ev) { this->message_port_pub(pmt::mp("events"), pmt::from_long(ev)); });  // 3 -> loss of lock (:1023)
}

// :617-641: an int event from the telemetry decoder; 1 = telemetry fault -> force the loss-of-lock condition
void b200_tracking_gr::msg_handler_telemetry_to_trk(const pmt::pmt_t& msg)
{
#ifdef GB200_WITH_GNURADIO
    try
        {
            if (pmt::any_ref(msg).type().hash_code() == typeid(int).hash_code()) d_impl->telemetry_to_trk(boost::any_cast<int>(pmt::any_ref(msg)));
        }
    catch (const std::exception&)
        {
        }
#else
    if (pmt::is_integer(msg)) d_impl->telemetry_to_trk(static_cast<int>(pmt::to_long(msg)));
#endif
}

// :607-614
void b200_tracking_gr::forecast(int noutput_items, gr_vector_int& ninput_items_required)
{
    if (noutput_items != 0) ninput_items_required[0] = static_cast<int32_t>(d_vector_length) * 2;
}

// :1728-2094 through the block object: plan -> B200_Multicorrelator_Real_Codes::Carrier_wipeoff_multicorrelator_resampler -> finish
int b200_tracking_gr::general_work(int noutput_items, gr_vector_int& ninput_items, gr_vector_const_void_star& input_items, gr_vector_void_star& output_items)
{
    (void)noutput_items;
    auto* out = reinterpret_cast<Gnss_Synchro*>(output_items[0]);
    int consumed = 0;
    const int produced = d_impl->general_work(input_items[0], ninput_items[0], this->nitems_read(0), out, &consumed);
    this->consume_each(consumed);
    return produced;
}

// ------------------------------------------------------------------ acquisition (pcps_acquisition.cc:50-52,902-1031)
b200_acquisition_gr::b200_acquisition_gr(std::shared_ptr<b200_pcps_acquisition> impl, size_t item_size)
    : gr::block("b200_pcps_acquisition", gr::io_signature::make(1, 1, static_cast<int>(item_size)), gr::io_signature::make(0, 1, sizeof(Gnss_Synchro))), d_impl(std::move(impl))
{
    this->message_port_register_out(pmt::mp("events"));
    d_impl->set_event_sink([this](int ev) { this->message_port_pub(pmt::mp("events"), pmt::from_long(ev)); });  // 1 positive / 2 negative (:362,389)
}

int b200_acquisition_gr::general_work(int noutput_items, gr_vector_int& ninput_items, gr_vector_const_void_star& input_items, gr_vector_void_star& output_items)
{
    (void)noutput_items;
    (void)output_items;
    this->consume_each(d_impl->general_work(input_items[0], ninput_items[0]));  // states 0 / 1 / 2: the FSM says how much it took
    return 0;
}

// ------------------------------------------------------------------ the other acquisition strategies
b200_fine_doppler_acquisition_gr::b200_fine_doppler_acquisition_gr(std::shared_ptr<b200_pcps_acquisition_fine_doppler> impl)
    : gr::block("b200_pcps_acquisition_fine_doppler", gr::io_signature::make(1, 1, sizeof(gr_complex)), gr::io_signature::make(0, 1, sizeof(Gnss_Synchro))), d_impl(std::move(impl))
{
    this->message_port_register_out(pmt::mp("events"));
    d_impl->set_event_sink([this](int ev) { this->message_port_pub(pmt::mp("events"), pmt::from_long(ev)); });
}
void b200_fine_doppler_acquisition_gr::forecast(int noutput_items, gr_vector_int& ninput_items_required)
{
    if (noutput_items != 0) ninput_items_required[0] = d_impl->fft_size();  // :171-178
}
int b200_fine_doppler_acquisition_gr::general_work(int noutput_items, gr_vector_int& ninput_items, gr_vector_const_void_star& input_items, gr_vector_void_star& output_items)
{
    (void)output_items;
    this->consume_each(d_impl->general_work(noutput_items, reinterpret_cast<const std::complex<float>*>(input_items[0]), ninput_items[0]));
    return 0;
}

b200_e5a_caf_acquisition_gr::b200_e5a_caf_acquisition_gr(std::shared_ptr<b200_e5a_noncoherent_iq_acquisition_caf> impl)
    : gr::block("b200_e5a_noncoherent_iq_acquisition_caf", gr::io_signature::make(1, 1, sizeof(gr_complex)), gr::io_signature::make(0, 1, sizeof(Gnss_Synchro))), d_impl(std::move(impl))
{
    this->message_port_register_out(pmt::mp("events"));
    d_impl->set_event_sink([this](int ev) { this->message_port_pub(pmt::mp("events"), pmt::from_long(ev)); });
}
int b200_e5a_caf_acquisition_gr::general_work(int noutput_items, gr_vector_int& ninput_items, gr_vector_const_void_star& input_items, gr_vector_void_star& output_items)
{
    (void)noutput_items;
    (void)output_items;
    this->consume_each(d_impl->general_work(reinterpret_cast<const std::complex<float>*>(input_items[0]), ninput_items[0]));
    return 0;
}

b200_quicksync_acquisition_gr::b200_quicksync_acquisition_gr(std::shared_ptr<b200_pcps_quicksync_acquisition> impl)
    : gr::block("b200_pcps_quicksync_acquisition", gr::io_signature::make(1, 1, static_cast<int>(sizeof(gr_complex)) * impl->item_length()), gr::io_signature::make(0, 1, sizeof(Gnss_Synchro))),
      d_impl(std::move(impl))
{
    this->message_port_register_out(pmt::mp("events"));
    d_impl->set_event_sink([this](int ev) { this->message_port_pub(pmt::mp("events"), pmt::from_long(ev)); });
}
int b200_quicksync_acquisition_gr::general_work(int noutput_items, gr_vector_int& ninput_items, gr_vector_const_void_star& input_items, gr_vector_void_star& output_items)
{
    (void)noutput_items;
    (void)output_items;
    this->consume_each(d_impl->general_work(reinterpret_cast<const std::complex<float>*>(input_items[0]), ninput_items[0]));  // vector items
    return 0;
}

std::shared_ptr<b200_tracking_gr> make_b200_tracking_gr(std::shared_ptr<b200_dll_pll_veml_tracking> impl, size_t item_size, unsigned vector_length)
{
    return std::make_shared<b200_tracking_gr>(std::move(impl), item_size, vector_length);
}
std::shared_ptr<b200_acquisition_gr> make_b200_acquisition_gr(std::shared_ptr<b200_pcps_acquisition> impl, size_t item_size)
{
    return std::make_shared<b200_acquisition_gr>(std::move(impl), item_size);
}
}  // namespace gb200h
